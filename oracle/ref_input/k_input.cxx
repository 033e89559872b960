/* TEST INFRASTRUCTURE ONLY.  The reference's .ini reader (src/input.cxx, Boost replaced by oracle/shim) compiled in place
 * as its own small library (oracle/_ref/liboracle_input.so — the main harness carries an in-memory Input stand-in with the
 * same member names, so the real one cannot live in the same library).  Values come back as text, errors as the text of the
 * std::runtime_error the reference throws. */
#include <cstdio>
#include <cstring>
#include <sstream>
#include <iomanip>
#include <stdexcept>
#include <string>
#include <vector>

#include "master.h"
#include "input.h"
#define API extern "C" __attribute__((visibility("default")))

namespace
{
    struct Handle { Master master; Input* input = nullptr; };
    void put(char* dst, int n, const std::string& s) { std::snprintf(dst, n, "%s", s.c_str()); }
    template<typename T> std::string show(const T& v) { std::ostringstream ss; ss << std::setprecision(17) << v; return ss.str(); }
    template<> std::string show<bool>(const bool& v) { return v ? "true" : "false"; }
    template<typename T> T parse(const char* s) { std::istringstream ss(s); T v{}; ss >> v; return v; }
    template<> std::string parse<std::string>(const char* s) { return s; }
    template<> bool parse<bool>(const char* s) { return std::strcmp(s, "true") == 0 || std::strcmp(s, "1") == 0; }

    template<typename T>
    int get(Handle* h, const char* b, const char* i, const char* s, int has_def, const char* def, char* out, int n)
    {
        try
        {
            const T v = has_def ? h->input->get_item<T>(b, i, s, parse<T>(def)) : h->input->get_item<T>(b, i, s);
            put(out, n, show<T>(v));
            return 0;
        }
        catch (std::exception& e) { put(out, n, e.what()); return 1; }
    }
    template<typename T>
    int get_list(Handle* h, const char* b, const char* i, const char* s, char* out, int n)
    {
        try
        {
            const std::vector<T> v = h->input->get_list<T>(b, i, s);
            std::string r;
            for (std::size_t q = 0; q < v.size(); ++q) r += (q ? "," : "") + show<T>(v[q]);
            put(out, n, r);
            return 0;
        }
        catch (std::exception& e) { put(out, n, e.what()); return 1; }
    }
}

// returns NULL and the exception text in err when the constructor throws
API void* refin_open(const char* path, char* err, int errlen)
{
    Handle* h = new Handle;
    try { h->master.start(); h->input = new Input(h->master, path); }
    catch (std::exception& e) { put(err, errlen, e.what()); delete h; return nullptr; }
    return h;
}
API void refin_close(void* p) { Handle* h = static_cast<Handle*>(p); delete h->input; delete h; }

// type: 'i' int, 'd' double, 'b' bool, 's' string; returns 0 and the value as text, or 1 and the error text
API int refin_get(void* p, char type, const char* b, const char* i, const char* s, int has_def, const char* def, char* out, int n)
{
    Handle* h = static_cast<Handle*>(p);
    switch (type)
    {
        case 'i': return get<int>(h, b, i, s, has_def, def, out, n);
        case 'd': return get<double>(h, b, i, s, has_def, def, out, n);
        case 'b': return get<bool>(h, b, i, s, has_def, def, out, n);
        default:  return get<std::string>(h, b, i, s, has_def, def, out, n);
    }
}
API int refin_get_list(void* p, char type, const char* b, const char* i, const char* s, char* out, int n)
{
    Handle* h = static_cast<Handle*>(p);
    switch (type)
    {
        case 'i': return get_list<int>(h, b, i, s, out, n);
        case 'd': return get_list<double>(h, b, i, s, out, n);
        case 'b': return get_list<bool>(h, b, i, s, out, n);
        default:  return get_list<std::string>(h, b, i, s, out, n);
    }
}
