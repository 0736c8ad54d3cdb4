// oracle/compat/ccutils/mpi/mpi_macros.hpp — TEST INFRASTRUCTURE (oracle).
// Re-creation of the ccutils print / section / JSON macros from their call sites in the
// reference (dp.cpp:174,180,248,275-295; fsdp.cpp:192,199,229,268,372,406-444) and from
// what plots/parser.py expects back (section id -> global json_data + per-rank JSON under
// "ccutils_rank_json"). The exact ccutils text grammar is not available; this one is:
//
//   #CCUTILS_SECTION_BEGIN <id> "<title>"
//   #CCUTILS_GLOBAL_JSON <id> {...}                (rank 0)
//   #CCUTILS_RANK_JSON <id> <rank> {...}           (every rank, in rank order)
//   #CCUTILS_SECTION_END <id>
#pragma once
#include <mpi.h>
#include <stdio.h>

#include <iostream>
#include <sstream>
#include <string>
#include <type_traits>
#include <vector>

namespace ccutils_compat {
inline int &rank() { static int r = 0; return r; }
inline int &size() { static int s = 1; return s; }

inline std::string json_escape(const std::string &s) {
    std::string o;
    for (char c : s) {
        if (c == '"' || c == '\\') { o += '\\'; o += c; }
        else if (c == '\n') o += "\\n";
        else o += c;
    }
    return o;
}
inline std::string to_json(const std::string &v) { return "\"" + json_escape(v) + "\""; }
inline std::string to_json(const char *v) { return to_json(std::string(v)); }
inline std::string to_json(char *v) { return to_json(std::string(v)); }
inline std::string to_json(bool v) { return v ? "true" : "false"; }
template <typename T, typename std::enable_if<std::is_arithmetic<T>::value, int>::type = 0>
inline std::string to_json(T v) {
    std::ostringstream os;
    os.precision(9);
    os << v;
    return os.str();
}
template <typename T> inline std::string to_json(const std::vector<T> &v) {
    std::string o = "[";
    for (size_t i = 0; i < v.size(); ++i) o += (i ? "," : "") + to_json(v[i]);
    return o + "]";
}
struct JsonObj {
    std::string body;
    template <typename T> void put(const char *k, const T &v) {
        if (!body.empty()) body += ",";
        body += "\"" + std::string(k) + "\":" + to_json(v);
    }
    std::string str() const { return "{" + body + "}"; }
};
// every rank prints `text` in rank order (token passing through barriers)
inline void ordered_print(const std::string &text) {
    for (int r = 0; r < size(); ++r) {
        if (r == rank()) { fputs(text.c_str(), stdout); fflush(stdout); }
        MPI_Barrier(MPI_COMM_WORLD);
    }
}
} // namespace ccutils_compat

#define CCUTILS_MPI_INIT                                                                           \
    MPI_Comm_rank(MPI_COMM_WORLD, &ccutils_compat::rank());                                        \
    MPI_Comm_size(MPI_COMM_WORLD, &ccutils_compat::size());

#define CCUTILS_MPI_PRINT_ONCE(...)                                                                \
    do { if (ccutils_compat::rank() == 0) { __VA_ARGS__; } } while (0);

#define CCUTILS_MPI_ALL_PRINT(...)                                                                 \
    do {                                                                                           \
        char *__cc_buf = nullptr; size_t __cc_len = 0;                                             \
        FILE *fp = open_memstream(&__cc_buf, &__cc_len);                                           \
        { __VA_ARGS__; }                                                                           \
        fclose(fp);                                                                                \
        ccutils_compat::ordered_print(std::string(__cc_buf, __cc_len));                            \
        free(__cc_buf);                                                                            \
    } while (0);

#define CCUTILS_MPI_SECTION_DEF(id, title)                                                         \
    ccutils_compat::JsonObj __cc_global_##id, __cc_rank_##id;                                      \
    const char *__cc_title_##id = title;

#define CCUTILS_MPI_GLOBAL_JSON_PUT(id, key, val) __cc_global_##id.put(key, val);
#define CCUTILS_SECTION_JSON_PUT(id, key, val) __cc_rank_##id.put(key, val);

#define CCUTILS_MPI_SECTION_END(id)                                                                \
    do {                                                                                           \
        if (ccutils_compat::rank() == 0) {                                                         \
            printf("#CCUTILS_SECTION_BEGIN %s \"%s\"\n", #id, __cc_title_##id);                    \
            printf("#CCUTILS_GLOBAL_JSON %s %s\n", #id, __cc_global_##id.str().c_str());           \
            fflush(stdout);                                                                        \
        }                                                                                          \
        MPI_Barrier(MPI_COMM_WORLD);                                                               \
        ccutils_compat::ordered_print("#CCUTILS_RANK_JSON " + std::string(#id) + " " +             \
                                      std::to_string(ccutils_compat::rank()) + " " +               \
                                      __cc_rank_##id.str() + "\n");                                \
        if (ccutils_compat::rank() == 0) { printf("#CCUTILS_SECTION_END %s\n", #id); fflush(stdout); } \
        MPI_Barrier(MPI_COMM_WORLD);                                                               \
    } while (0);
