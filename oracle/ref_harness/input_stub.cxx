/* TEST INFRASTRUCTURE ONLY (oracle/_ref harness).
 *
 * The reference's Input class (include/input.h:36-52, src/input.cxx) parses .ini
 * files with Boost, which is not in this image.  This file supplies the same
 * members backed by an in-memory "block.item[sub]" -> string table, so that the
 * reference's own constructors (Grid, Master::init, Advec_2i5, ...) can run.
 */
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "input.h"
#include "master.h"

std::map<std::string, std::string>& mhh_oracle_items()
{
    static std::map<std::string, std::string> items;
    return items;
}

namespace
{
    std::string make_key(const std::string& b, const std::string& i, const std::string& s)
    {
        return s.empty() ? b + "." + i : b + "." + i + "[" + s + "]";
    }

    template<typename T> T convert(const std::string& v)
    {
        std::istringstream ss(v);
        T out;
        ss >> out;
        return out;
    }
    template<> std::string convert<std::string>(const std::string& v) { return v; }
    template<> bool convert<bool>(const std::string& v) { return v == "1" || v == "true"; }

    bool lookup(const std::string& b, const std::string& i, const std::string& s, std::string& out)
    {
        auto& items = mhh_oracle_items();
        auto it = items.find(make_key(b, i, s));
        if (it == items.end() && !s.empty())
            it = items.find(make_key(b, i, ""));
        if (it == items.end())
            return false;
        out = it->second;
        return true;
    }
}

Input::Input(Master& masterin, const std::string&) : master(masterin) {}

template<typename T>
T Input::get_item(const std::string& b, const std::string& i, const std::string& s)
{
    std::string v;
    if (!lookup(b, i, s, v))
        throw std::runtime_error("oracle input stub: missing item " + make_key(b, i, s));
    return convert<T>(v);
}

template<typename T>
T Input::get_item(const std::string& b, const std::string& i, const std::string& s, const T def)
{
    std::string v;
    if (!lookup(b, i, s, v))
        return def;
    return convert<T>(v);
}

template<typename T>
std::vector<T> Input::get_list(const std::string& b, const std::string& i, const std::string& s)
{
    std::string v;
    if (!lookup(b, i, s, v))
        throw std::runtime_error("oracle input stub: missing list " + make_key(b, i, s));
    std::vector<T> out;
    std::istringstream ss(v);
    std::string tok;
    while (std::getline(ss, tok, ','))
        if (!tok.empty())
            out.push_back(convert<T>(tok));
    return out;
}

template<typename T>
std::vector<T> Input::get_list(const std::string& b, const std::string& i, const std::string& s, const std::vector<T> def)
{
    std::string v;
    if (!lookup(b, i, s, v))
        return def;
    return get_list<T>(b, i, s);
}

void Input::print_unused_items() {}
void Input::flag_as_used(const std::string&, const std::string&, const std::string&) {}

#define INSTANTIATE(T) \
    template T Input::get_item<T>(const std::string&, const std::string&, const std::string&); \
    template T Input::get_item<T>(const std::string&, const std::string&, const std::string&, const T); \
    template std::vector<T> Input::get_list<T>(const std::string&, const std::string&, const std::string&); \
    template std::vector<T> Input::get_list<T>(const std::string&, const std::string&, const std::string&, const std::vector<T>);

INSTANTIATE(int)
INSTANTIATE(double)
INSTANTIATE(float)
INSTANTIATE(bool)
INSTANTIATE(std::string)
INSTANTIATE(unsigned long)
