// cpp/b200/host_runtime.hpp — host-side support shared by the b200 proxy drivers.
//
// The reference leans on two un-vendored packages for everything around its hot loop:
// MPI (process group, barriers, string gathers) and ccutils (host timers, rank-ordered
// printing, JSON "sections", the easyargs CLI generator) — see dp.cpp:9,28-30,124 and
// fsdp.cpp:12,22-23,182. The b200 build is self-contained: the process group comes from
// the C ABI's rendezvous (dlnb_host_barrier / dlnb_host_allgather) and the small pieces
// below replace the ccutils subset. The stdout grammar of a section is
//
//   #CCUTILS_SECTION_BEGIN <id> "<title>"
//   #CCUTILS_GLOBAL_JSON <id> {...}                 rank 0
//   #CCUTILS_RANK_JSON <id> <rank> {...}            every rank, in rank order
//   #CCUTILS_SECTION_END <id>
//
// with the keys of the reference's sections (dp.cpp:275-295, fsdp.cpp:406-444), which is
// what plots/parser.py consumes (global json_data + per-rank "ccutils_rank_json").
#pragma once

#include <dlnb.h>
#include <signal.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <sstream>
#include <string>
#include <type_traits>
#include <vector>

namespace b200 {

// ---- fail-fast, like CCUTILS_GPU_CHECK / MPI_ERRORS_ARE_FATAL --------------------------
[[noreturn]] inline void die(const std::string &what) {
    fprintf(stderr, "[rank %d] %s\n", dlnb_rank(), what.c_str());
    fflush(stderr);
    _exit(1);
}
inline void check(int rc, const char *what) {
    if (rc != 0) die(std::string(what) + ": " + dlnb_last_error());
}

// ---- process group ------------------------------------------------------------------------
struct World {
    int rank = 0, size = 1, local_rank = 0;
    std::string session;
};
inline int env_int(const char *a, const char *b, int dflt) {
    const char *v = getenv(a);
    if (!v && b) v = getenv(b);
    return v ? atoi(v) : dflt;
}
// rank / size come from the launcher (tools/dlnb_launch sets DLNB_*; torchrun's names work too)
inline World world_from_env() {
    World w;
    w.rank = env_int("DLNB_RANK", "RANK", 0);
    w.size = env_int("DLNB_WORLD", "WORLD_SIZE", 1);
    w.local_rank = env_int("DLNB_LOCAL_RANK", "LOCAL_RANK", w.rank);
    const char *s = getenv("DLNB_SESSION");
    if (s) w.session = s;
    else if (const char *p = getenv("MASTER_PORT")) w.session = std::string("port-") + p;
    else w.session = "solo-" + std::to_string(getpid());
    return w;
}
inline void host_barrier() { check(dlnb_host_barrier(), "host barrier"); }
template <typename T> inline std::vector<T> host_allgather(const T &mine) {
    static_assert(std::is_trivially_copyable<T>::value, "POD only");
    std::vector<T> all((size_t)dlnb_world());
    check(dlnb_host_allgather(&mine, sizeof(T), all.data()), "host allgather");
    return all;
}
inline std::vector<std::string> host_allgather_strings(const std::string &mine, size_t width = 256) {
    std::vector<char> me(width, 0), all(width * (size_t)dlnb_world(), 0);
    strncpy(me.data(), mine.c_str(), width - 1);
    check(dlnb_host_allgather(me.data(), width, all.data()), "host allgather");
    std::vector<std::string> out;
    for (int r = 0; r < dlnb_world(); ++r) out.emplace_back(all.data() + (size_t)r * width);
    return out;
}

// ---- wall-clock timers: one float of SECONDS per start/stop pair ---------------------------
inline double wtime() {
    timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec;
}
struct HostTimer {
    std::vector<float> vals;
    double t0 = 0.0;
    void start() { t0 = wtime(); }
    void stop() { vals.push_back((float)(wtime() - t0)); }
    void clear() { vals.clear(); }
};

// ---- rank-ordered printing ----------------------------------------------------------------------
inline void print_once(const std::string &text, FILE *to = stdout) {
    if (dlnb_rank() == 0) { fputs(text.c_str(), to); fflush(to); }
}
inline void print_all_ordered(const std::string &text) {
    for (int r = 0; r < dlnb_world(); ++r) {
        if (r == dlnb_rank()) { fputs(text.c_str(), stdout); fflush(stdout); }
        host_barrier();
    }
}

// ---- JSON sections ---------------------------------------------------------------------------------
class Json {
  public:
    template <typename T> Json &put(const std::string &key, const T &v) {
        if (!body_.empty()) body_ += ",";
        body_ += quote(key) + ":" + enc(v);
        return *this;
    }
    std::string str() const { return "{" + body_ + "}"; }

  private:
    std::string body_;
    static std::string quote(const std::string &s) {
        std::string o = "\"";
        for (char c : s) {
            if (c == '"' || c == '\\') o += '\\';
            if (c == '\n') { o += "\\n"; continue; }
            o += c;
        }
        return o + "\"";
    }
    static std::string enc(const std::string &v) { return quote(v); }
    static std::string enc(const char *v) { return quote(v); }
    static std::string enc(bool v) { return v ? "true" : "false"; }
    template <typename T, typename std::enable_if<std::is_arithmetic<T>::value, int>::type = 0>
    static std::string enc(T v) {
        std::ostringstream os;
        os.precision(9);
        os << v;
        return os.str();
    }
    template <typename T> static std::string enc(const std::vector<T> &v) {
        std::string o = "[";
        for (size_t i = 0; i < v.size(); ++i) o += (i ? "," : "") + enc(v[i]);
        return o + "]";
    }
};

struct Section {
    std::string id, title;
    Json global, per_rank;
    Section(std::string i, std::string t) : id(std::move(i)), title(std::move(t)) {}
    void emit() {
        if (dlnb_rank() == 0) {
            printf("#CCUTILS_SECTION_BEGIN %s \"%s\"\n#CCUTILS_GLOBAL_JSON %s %s\n", id.c_str(), title.c_str(),
                   id.c_str(), global.str().c_str());
            fflush(stdout);
        }
        host_barrier();
        print_all_ordered("#CCUTILS_RANK_JSON " + id + " " + std::to_string(dlnb_rank()) + " " + per_rank.str() + "\n");
        if (dlnb_rank() == 0) { printf("#CCUTILS_SECTION_END %s\n", id.c_str()); fflush(stdout); }
        host_barrier();
    }
};

// ---- CLI: positional required arguments, then -w/-r/-d/-m/-h (dp.cpp:108-122, fsdp.cpp:165-180);
//      additive: -t <fp32|fp16|bf16>, -a <auto|oneshot|twoshot|multimem>, -c <ctas>, -s <spin ctas>, -g ---
struct Cli {
    std::vector<std::string> positional;
    int warmup = 3, runs = 5, min_exectime = 0, ctas = 0, spin_ctas = -1 /* library default: one per SM */, spin_threads = 0;
    std::string devices, dtype = "fp16", algo = "auto", spin_mode;
    bool help = false, ok = true, graph = false;
};
inline Cli parse_cli(int argc, char **argv) {
    Cli c;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto value = [&](std::string &dst) { if (i + 1 < argc) dst = argv[++i]; else c.ok = false; };
        auto ivalue = [&](int &dst) { std::string s; value(s); if (c.ok) dst = atoi(s.c_str()); };
        if (a == "-h") c.help = true;
        else if (a == "-g") c.graph = true;
        else if (a == "-w") ivalue(c.warmup);
        else if (a == "-r") ivalue(c.runs);
        else if (a == "-m") ivalue(c.min_exectime);
        else if (a == "-d") value(c.devices);
        else if (a == "-t") value(c.dtype);
        else if (a == "-a") value(c.algo);
        else if (a == "-c") ivalue(c.ctas);
        else if (a == "-s") ivalue(c.spin_ctas);
        else if (a == "-S") value(c.spin_mode);
        else if (a == "-T") ivalue(c.spin_threads);
        else if (a.size() > 1 && a[0] == '-' && !isdigit((unsigned char)a[1])) c.ok = false;
        else c.positional.push_back(a);
    }
    return c;
}
inline void print_options_help() {
    printf("Options:\n"
           "  -w <warmups>       Number of warm-up iterations (default: 3)\n"
           "  -r <runs>          Number of iterations to run (default: 5)\n"
           "  -d <devices>       Comma-separated list of devices (default: \"\")\n"
           "  -m <min_exectime>  Minimum total execution time in seconds (overrides runs) (default: 0)\n"
           "  -t <dtype>         fp32 | fp16 | bf16 (default: fp16)          [b200 addition]\n"
           "  -a <algo>          auto | oneshot | twoshot | multimem allreduce (default: auto)  [b200 addition]\n"
           "  -c <ctas>          CTAs per collective kernel, 0 = auto        [b200 addition]\n"
           "  -s <spin_ctas>     CTAs of the compute-emulation kernel (default: one per SM)  [b200 addition]\n"
           "  -T <spin_threads>  Threads per CTA of the compute-emulation kernel (default: 256)  [b200 addition]\n"
           "  -S <spin_mode>     What the emulated compute does: sleep | fma | hbm (default: fma)  [b200 addition]\n"
           "  -g                 Capture one iteration as a CUDA graph and replay it  [b200 addition]\n"
           "  -h                 Show help\n");
}
inline int dtype_from_name(const std::string &s) {
    if (s == "fp32" || s == "f32" || s == "float32") return DLNB_F32;
    if (s == "fp16" || s == "f16" || s == "float16") return DLNB_F16;
    if (s == "bf16" || s == "bfloat16") return DLNB_BF16;
    return -1;
}
inline int algo_from_name(const std::string &s) {
    if (s == "auto") return DLNB_ALGO_AUTO;
    if (s == "oneshot") return DLNB_ALGO_ONESHOT;
    if (s == "twoshot") return DLNB_ALGO_TWOSHOT;
    if (s == "multimem") return DLNB_ALGO_MULTIMEM;
    return -1;
}
inline size_t dtype_bytes(int dt) { return dt == DLNB_F32 ? 4 : 2; }
// footprint of the emulated compute from the command line (-s / -T / -S); call after dlnb_init
inline void apply_spin_cli(const Cli &c) {
    if (c.spin_ctas >= 0) check(dlnb_set("spin_ctas", c.spin_ctas), "dlnb_set spin_ctas");
    if (c.spin_threads > 0) check(dlnb_set("spin_threads", c.spin_threads), "dlnb_set spin_threads");
    if (!c.spin_mode.empty()) {
        const long m = c.spin_mode == "sleep" ? 0 : c.spin_mode == "fma" ? 1 : c.spin_mode == "hbm" ? 2 : -1;
        if (m < 0) die("unknown spin mode '" + c.spin_mode + "' (sleep | fma | hbm)");
        check(dlnb_set("spin_mode", m), "dlnb_set spin_mode");
    }
}
inline const char *spin_mode_name() {
    const long m = dlnb_get("spin_mode");
    return m == 0 ? "sleep" : m == 1 ? "fma" : "hbm";
}

// ---- *_loop binaries: run forever and print nothing (dp.cpp:251-255), unless
//      DLNB_LOOP_ITERS bounds the loop or SIGTERM/SIGINT asks for a summary ---------------------------
inline volatile sig_atomic_t &stop_flag() { static volatile sig_atomic_t f = 0; return f; }
inline void on_stop_signal(int) { stop_flag() = 1; }
inline void install_stop_handlers() {
    struct sigaction sa;
    memset(&sa, 0, sizeof(sa));
    sa.sa_handler = on_stop_signal;
    sigaction(SIGTERM, &sa, nullptr);
    sigaction(SIGINT, &sa, nullptr);
}
inline long loop_iters_from_env() {
    const char *v = getenv("DLNB_LOOP_ITERS");
    return v ? atol(v) : -1; // -1: forever
}

// One proxy iteration, either issued call by call or captured once and replayed (-g).
template <typename Fn> struct IterationRunner {
    Fn body;
    dlnb_graph_t graph = nullptr;
    bool use_graph;
    IterationRunner(Fn f, bool g) : body(f), use_graph(g) {}
    void operator()() {
        if (!use_graph) { body(); return; }
        if (!graph) {
            check(dlnb_graph_begin(), "graph begin");
            body();
            check(dlnb_graph_end(&graph), "graph end");
        }
        check(dlnb_graph_launch(graph), "graph launch");
        check(dlnb_compute_sync(), "graph sync");
    }
};
template <typename Fn> inline IterationRunner<Fn> make_runner(Fn f, bool g) { return IterationRunner<Fn>(f, g); }

inline float median(std::vector<float> v) {
    if (v.empty()) return 0.f;
    std::sort(v.begin(), v.end());
    return v[v.size() / 2];
}
// element-wise max over ranks of a per-iteration series (device-timed iteration = max over ranks)
inline std::vector<float> max_over_ranks(const std::vector<float> &mine) {
    const size_t n = mine.size();
    std::vector<float> all(n * (size_t)dlnb_world());
    if (n) check(dlnb_host_allgather(mine.data(), n * sizeof(float), all.data()), "host allgather");
    std::vector<float> out(n, 0.f);
    for (int r = 0; r < dlnb_world(); ++r)
        for (size_t i = 0; i < n; ++i) out[i] = std::max(out[i], all[(size_t)r * n + i]);
    return out;
}

} // namespace b200
