// dp / dp_loop — data-parallel training proxy, b200 build.
//
// Same CLI, inputs and output section as the reference's cpp/data_parallel/dp.cpp
// (`dp <model> <num_buckets> <base_path> [-w][-r][-d][-m][-h]`, model_stats/<model>.txt,
// section "dp" with the keys of dp.cpp:275-295); dp_loop is this file with -DPROXY_LOOP.
// What differs is underneath: the forward/backward "compute" is a calibrated %globaltimer
// spin kernel on the GPU's compute stream instead of a host usleep (dp.cpp:93,98), each
// bucket's Allreduce is a hand-written sm_100a kernel chained to the end of its backward
// slice by a device event, and the host only blocks in WaitAll.
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <vector>

#include "../data_types.hpp"
#include "../netcommunicators.hpp"
#include "../proxy_classes.hpp"
#include "../utils.hpp"

constexpr Device device = Device::GPU;

static b200::HostTimer timer_runtime, timer_barrier;
static bool blocking_compute = false; // DLNB_BLOCKING_COMPUTE=1: host waits for each spin, like usleep

static inline void emulate_compute(uint64_t usec) {
    b200::check(dlnb_compute_spin((double)usec), "compute spin");
    if (blocking_compute) b200::check(dlnb_compute_sync(), "compute sync");
}

// One iteration: forward of the whole model, then per bucket a slice of backward followed by
// the bucket's non-blocking Allreduce; the exposed communication tail is timed as "barrier".
template <typename T>
int run_data_parallel(Tensor<T, device> **grad_ptrs, Tensor<T, device> **sum_grad_ptrs, int num_buckets,
                      uint64_t *params_per_bucket, uint64_t fwd_rt_whole_model, float bwd_rt_per_B,
                      ProxyCommunicator *communicator) {
    emulate_compute(fwd_rt_whole_model);
    for (int i = 0; i < num_buckets; i++) {
        emulate_compute((uint64_t)bwd_rt_per_B); // usleep(float) truncates the same way
        proxy_iallreduce(communicator, grad_ptrs[i]->data, sum_grad_ptrs[i]->data, params_per_bucket[i], i); // int when it fits
    }
    timer_barrier.start();
    communicator->WaitAll(num_buckets);
    timer_barrier.stop();
    return 0;
}

static void print_help(const char *prog) {
    printf("Usage: %s <model> <num_buckets> <base_path> [options]\n\n"
           "Arguments:\n"
           "  model        Name of the model to use (model_stats/<model>.txt)\n"
           "  num_buckets  Number of parameter buckets\n"
           "  base_path    Base path for the repository (relative to the current directory)\n\n",
           prog);
    b200::print_options_help();
}

template <typename T> static int proxy_main(const b200::Cli &cli, const b200::World &w) {
    const std::string model_name = cli.positional[0];
    const int num_buckets = atoi(cli.positional[1].c_str());
    if (num_buckets <= 0) {
        fprintf(stderr, "Error: num_buckets must be positive\n");
        return 1;
    }
    const fs::path stats_file = get_dnnproxy_base_path(cli.positional[2].c_str()) / "model_stats" / (model_name + ".txt");
    if (!fs::exists(stats_file)) {
        std::cerr << "Error: model stats file does not exist: " << stats_file << "\n";
        return -1;
    }
    std::map<std::string, uint64_t> stats = get_model_stats(stats_file.string());
    const uint64_t fwd_rt_whole_model = stats["avgForwardTime"];                     // us
    const float bwd_rt_per_B = (float)(stats["avgBackwardTime"] / (uint64_t)num_buckets); // us, integer division first
    const unsigned local_batch_size = (unsigned)stats["batchSize"];
    const uint64_t total_model_size = stats["modelSize"];

    std::vector<uint64_t> params_per_bucket((size_t)num_buckets);
    for (int i = 0; i < num_buckets; i++)
        params_per_bucket[i] = total_model_size / num_buckets + ((uint64_t)i < total_model_size % num_buckets ? 1 : 0);
    // (buckets beyond 2^31-1 elements, where the reference's int count wraps, take the backend's 64-bit entry)

    const int my_device = set_local_device(w.local_rank, cli.devices.c_str());
    b200::check(dlnb_init(w.rank, w.size, my_device, w.session.c_str()), "dlnb_init");
    print_topology_graph();
    {
        char line[128];
        snprintf(line, sizeof(line), "Rank %d using device %d\n", w.rank, my_device);
        b200::print_all_ordered(line);
    }
    b200::apply_spin_cli(cli);

    std::vector<int> everyone((size_t)w.size);
    for (int r = 0; r < w.size; r++) everyone[r] = r;
    ProxyBackend *communicator = new ProxyBackend(everyone, num_buckets, ProxyDtype<T>::value);
    communicator->set("algo", b200::algo_from_name(cli.algo));
    if (cli.ctas > 0) communicator->set("ctas", cli.ctas);

    std::vector<Tensor<T, device> *> grad_ptrs((size_t)num_buckets), sum_grad_ptrs((size_t)num_buckets);
    for (int i = 0; i < num_buckets; i++) {
        grad_ptrs[i] = new Tensor<T, device>(params_per_bucket[i]);
        sum_grad_ptrs[i] = new Tensor<T, device>(params_per_bucket[i]);
        // integer-valued gradients: sums are exact in every dtype, and no NaN/denormal paths
        b200::check(dlnb_fill_pattern(grad_ptrs[i]->data, params_per_bucket[i], ProxyDtype<T>::value, DLNB_PAT_INT, 0, w.rank),
                    "fill");
    }
    double spin_overhead_us = 0, spin_slope = 1;
    b200::check(dlnb_spin_calibrate(&spin_overhead_us, &spin_slope), "spin calibration");
    b200::host_barrier();

    auto iteration = b200::make_runner(
        [&]() {
            run_data_parallel<T>(grad_ptrs.data(), sum_grad_ptrs.data(), num_buckets, params_per_bucket.data(),
                                 fwd_rt_whole_model, bwd_rt_per_B, communicator);
        },
        cli.graph);
    std::vector<float> warmup_times;
    for (int i = 0; i < cli.warmup; i++) {
        const double t0 = b200::wtime();
        iteration();
        warmup_times.push_back((float)(b200::wtime() - t0));
    }
    int runs = cli.runs;
    if (cli.min_exectime > 0) {
        runs = estimate_runs(warmup_times, (uint64_t)cli.min_exectime);
        b200::print_once("Estimated runs based on warm-up times to meet minimum execution time: " + std::to_string(runs) + "\n");
    }

    std::vector<float> device_ms;
#ifdef PROXY_LOOP
    // Runs until killed and prints nothing, like the reference's dp_loop. DLNB_LOOP_ITERS=<n>
    // (or SIGTERM / SIGINT) ends the loop and reports what was measured.
    b200::install_stop_handlers();
    const long bound = b200::loop_iters_from_env();
    for (long it = 0; (bound < 0 || it < bound) && !b200::stop_flag(); it++) {
        const bool keep = bound >= 0 || device_ms.size() < 4096;
        if (keep) { timer_runtime.start(); b200::check(dlnb_iter_begin(), "iter begin"); }
        iteration();
        if (keep) {
            float ms = 0;
            b200::check(dlnb_iter_end(&ms), "iter end");
            timer_runtime.stop();
            device_ms.push_back(ms);
        } else {
            timer_barrier.clear();
        }
    }
    if (bound < 0) {
        // stopped by a signal: ranks left the loop at different iterations, so no collective
        // reporting — each rank says what it saw and leaves
        fprintf(stderr, "[rank %d] dp_loop stopped after %zu timed iterations, median %.3f ms (device)\n", w.rank,
                device_ms.size(), b200::median(device_ms));
        _exit(0);
    }
#else
    timer_barrier.clear();
    for (int it = 0; it < runs; it++) {
        timer_runtime.start();
        b200::check(dlnb_iter_begin(), "iter begin");
        iteration();
        float ms = 0;
        b200::check(dlnb_iter_end(&ms), "iter end");
        timer_runtime.stop();
        device_ms.push_back(ms);
    }
#endif

    // device time of every bucket's Allreduce in the last iteration -> bus bandwidth
    double bus_sum = 0;
    float slowest_ms = 0;
    if (cli.graph) { // slots are not host-visible inside a replay: time one plain iteration for the report
        run_data_parallel<T>(grad_ptrs.data(), sum_grad_ptrs.data(), num_buckets, params_per_bucket.data(),
                             fwd_rt_whole_model, bwd_rt_per_B, communicator);
        timer_barrier.vals.pop_back();
    }
    for (int i = 0; i < num_buckets; i++) {
        const float ms = communicator->slot_ms(i);
        const double bytes = (double)params_per_bucket[i] * sizeof(T);
        const double wire = w.size > 1 ? 2.0 * (w.size - 1) / w.size * bytes : 2.0 * bytes;
        bus_sum += wire / (ms * 1e-3) / 1e9;
        slowest_ms = std::max(slowest_ms, ms);
    }

    char host_name[256] = "localhost";
    gethostname(host_name, sizeof(host_name) - 1);
    const std::pair<float, float> msg_stats = compute_msg_stats(params_per_bucket, 1);
    const uint64_t lower_bound_us = fwd_rt_whole_model + (uint64_t)num_buckets * (uint64_t)bwd_rt_per_B;
    const std::vector<float> device_ms_max = b200::max_over_ranks(device_ms);

    b200::Section dp("dp", "Data Parallelism");
    dp.global.put("model_name", model_name)
        .put("num_buckets", num_buckets)
        .put("local_batch_size", local_batch_size)
        .put("world_size", w.size)
        .put("fwd_rt_whole_model", fwd_rt_whole_model)
        .put("bwd_rt_per_bucket", bwd_rt_per_B)
        .put("total_model_size_params", total_model_size)
        .put("msg_size_avg_bytes", msg_stats.first * sizeof(T))
        .put("msg_size_std_bytes", msg_stats.second * sizeof(T))
        .put("device", device == Device::CPU ? "CPU" : "GPU")
        .put("backend", communicator->get_name())
        // b200 additions (never replacing the keys above)
        .put("dtype", ProxyDtype<T>::name)
        .put("arena", dlnb_arena_kind())
        .put("lower_bound_us", lower_bound_us)
        .put("device_iter_ms_max_over_ranks", device_ms_max)
        .put("device_iter_ms_median", b200::median(device_ms_max))
        .put("ratio_to_lower_bound", lower_bound_us ? b200::median(device_ms_max) * 1000.0 / (double)lower_bound_us : 0.0)
        .put("allreduce_busbw_GBps_avg", bus_sum / num_buckets)
        .put("allreduce_slowest_ms", slowest_ms)
        .put("spin_overhead_us", spin_overhead_us)
        .put("spin_slope", spin_slope)
        .put("spin_ctas", (long)dlnb_get("spin_ctas"))
        .put("spin_threads", (long)dlnb_get("spin_threads"))
        .put("spin_mode", b200::spin_mode_name())
        .put("spin_coresident", (long)dlnb_get("spin_coresident"))
        .put("cuda_graph", cli.graph)
        .put("kernel_launches", (uint64_t)dlnb_launch_count());
    dp.per_rank.put("runtimes", timer_runtime.vals)
        .put("barrier_time", timer_barrier.vals)
        .put("hostname", host_name)
        .put("device_runtimes_ms", device_ms);
    dp.emit();

    communicator->finalize();
    b200::check(dlnb_finalize(), "dlnb_finalize");
    return 0;
}

int main(int argc, char *argv[]) {
    const b200::Cli cli = b200::parse_cli(argc, argv);
    if (!cli.ok || cli.help || cli.positional.size() != 3) {
        print_help(argv[0]);
        return 1;
    }
    blocking_compute = getenv("DLNB_BLOCKING_COMPUTE") && atoi(getenv("DLNB_BLOCKING_COMPUTE")) != 0;
    const b200::World w = b200::world_from_env();
    switch (b200::dtype_from_name(cli.dtype)) {
    case DLNB_F32: return proxy_main<float>(cli, w);
    case DLNB_F16: return proxy_main<proxy_half>(cli, w);
    case DLNB_BF16: return proxy_main<proxy_bfloat16>(cli, w);
    }
    fprintf(stderr, "Error: unknown dtype '%s' (fp32 | fp16 | bf16)\n", cli.dtype.c_str());
    return 1;
}
