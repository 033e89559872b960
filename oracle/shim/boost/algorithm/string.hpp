/* TEST INFRASTRUCTURE ONLY — minimal stand-in for <boost/algorithm/string.hpp> (Boost is absent here): just what
 * src/fields.cxx:259 and src/input.cxx:101-146,373,410 use: boost::split(container, string, boost::is_any_of(..)
 * [, boost::token_compress_on]) and boost::trim(string) (both ends, std::isspace in the "C" locale). */
#pragma once
#include <string>
#include <vector>

namespace boost
{
    struct is_any_of { std::string set; explicit is_any_of(const char* s) : set(s) {} };
    enum token_compress_mode_type { token_compress_on, token_compress_off };

    template<class Container>
    Container& split(Container& out, const std::string& in, const is_any_of& pred, token_compress_mode_type mode = token_compress_off)
    {
        out.clear();
        std::string cur;
        bool last_was_sep = false;
        for (char ch : in)
        {
            if (pred.set.find(ch) != std::string::npos)
            {
                if (!(mode == token_compress_on && last_was_sep)) { out.push_back(cur); cur.clear(); }
                last_was_sep = true;
            }
            else { cur.push_back(ch); last_was_sep = false; }
        }
        out.push_back(cur);
        return out;
    }

    inline void to_lower(std::string& s)          // include/input_tools.h:74 (bool items: "TRUE" -> "true")
    {
        for (char& ch : s)
            if (ch >= 'A' && ch <= 'Z') ch = static_cast<char>(ch - 'A' + 'a');
    }

    inline void trim(std::string& s)
    {
        const char* ws = " \t\n\v\f\r";
        const std::size_t b = s.find_first_not_of(ws);
        if (b == std::string::npos) { s.clear(); return; }
        const std::size_t e = s.find_last_not_of(ws);
        s = s.substr(b, e - b + 1);
    }
}
