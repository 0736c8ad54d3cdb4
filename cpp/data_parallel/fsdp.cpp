// fsdp / fsdp_loop — fully-sharded data-parallel training proxy, b200 build.
//
// Same CLI, inputs and output section as the reference's cpp/data_parallel/fsdp.cpp
// (`fsdp <model> <num_units> <sharding_factor> <base_path> [-w][-r][-d][-m][-h]`, section
// "fsdp" with the keys of fsdp.cpp:406-444); fsdp_loop is this file with -DPROXY_LOOP.
// Per unit: prefetch Allgather of the next unit's shards while the current unit computes,
// blocking ReduceScatter of the gradients after backward, and — when sharding_factor <
// world_size (HSDP) — a non-blocking Allreduce of the shard across replicas. Compute is a
// %globaltimer spin kernel on the compute stream; collectives are sm_100a kernels over peer
// HBM on an ordered communication stream, so the ReduceScatter of unit u runs after the
// in-flight Allgather of unit u-1 into the shared gather buffer (the reference's NCCL build
// races them, fsdp.cpp:113-126 + proxy_classes.hpp:203,207).
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <vector>

#include "../data_types.hpp"
#include "../netcommunicators.hpp"
#include "../proxy_classes.hpp"
#include "../utils.hpp"

constexpr Device device = Device::GPU;

static b200::HostTimer timer_allgather, timer_reduce_scatter, timer_allgather_wait_fwd, timer_allgather_wait_bwd,
    timer_barrier, timer_runtime;
static bool blocking_compute = false;
static bool in_graph_capture = false;
static float first_allgather_ms = 0.f; // device time of the exposed Allgather of unit 0 (most recent iteration)
// what was LEFT of each prefetch Allgather after the compute it was hidden behind had ended (device-timed, ms): the
// reference's meaning of allgather_wait_fwd / _bwd, kept beside the host timers because compute is asynchronous here
static std::vector<float> exposed_wait_fwd_ms, exposed_wait_bwd_ms;
static float exposed_after_mark(ProxyCommunicator *comm, int slot) {
    float ms = 0.f;
    if (!in_graph_capture) b200::check(dlnb_slot_exposed(static_cast<ProxyBackend *>(comm)->handle(), slot, &ms), "slot exposed");
    return ms;
}

static inline void emulate_compute(float usec) {
    b200::check(dlnb_compute_spin((double)(uint64_t)usec), "compute spin"); // usleep(float) truncates
    if (blocking_compute) b200::check(dlnb_compute_sync(), "compute sync");
}

template <typename T>
void run_fsdp(Tensor<T, device> **shard_params, Tensor<T, device> *layer_params, Tensor<T, device> **allreduce_params,
              float fwd_rt_whole_unit, float bwd_rt_whole_unit, unsigned num_units, unsigned sharding_factor,
              uint64_t *max_params_per_shard, unsigned num_replicas, ProxyCommunicator *unit_comm,
              ProxyCommunicator *allreduce_comm) {
    (void)sharding_factor;
    auto shard_count = [&](unsigned u) { return static_cast<int>(max_params_per_shard[u]); };

    // parameters of the first unit: nothing to hide it behind
    timer_allgather.start();
    unit_comm->Allgather(shard_params[0]->data, shard_count(0), layer_params->data, shard_count(0));
    timer_allgather.stop();
    if (!in_graph_capture) first_allgather_ms = static_cast<ProxyBackend *>(unit_comm)->slot_ms(-1);

    // forward: gather unit u+1 while unit u computes (the last unit's forward is not emulated,
    // as in the reference: U-1 sleeps)
    for (unsigned u = 0; u + 1 < num_units; u++) {
        unit_comm->Iallgather(shard_params[u + 1]->data, shard_count(u + 1), layer_params->data, shard_count(u + 1), u + 1);
        emulate_compute(fwd_rt_whole_unit);
        b200::check(dlnb_compute_mark(), "compute mark");
        timer_allgather_wait_fwd.start();
        unit_comm->Wait(u + 1);
        timer_allgather_wait_fwd.stop();
        exposed_wait_fwd_ms.push_back(exposed_after_mark(unit_comm, (int)u + 1));
    }

    // backward, last unit first: gather unit u-1, compute unit u, scatter-reduce its gradients
    for (int u = (int)num_units - 1; u >= 0; u--) {
        if (u > 0)
            unit_comm->Iallgather(shard_params[u - 1]->data, shard_count(u - 1), layer_params->data, shard_count(u - 1), u - 1);
        emulate_compute(bwd_rt_whole_unit);
        b200::check(dlnb_compute_mark(), "compute mark");

        timer_reduce_scatter.start();
        unit_comm->Reduce_Scatter_block(layer_params->data, shard_params[u]->data, shard_count(u));
        timer_reduce_scatter.stop();

        if (num_replicas > 1)
            allreduce_comm->Iallreduce(shard_params[u]->data, allreduce_params[u]->data, shard_count(u), u);
        if (u > 0) {
            timer_allgather_wait_bwd.start();
            unit_comm->Wait(u - 1);
            timer_allgather_wait_bwd.stop();
            exposed_wait_bwd_ms.push_back(exposed_after_mark(unit_comm, u - 1));
        }
    }
    if (num_replicas > 1) {
        timer_barrier.start();
        allreduce_comm->WaitAll(num_units);
        timer_barrier.stop();
    }
}

static void print_help(const char *prog) {
    printf("Usage: %s <model> <num_units> <sharding_factor> <base_path> [options]\n\n"
           "Arguments:\n"
           "  model            Name of the model to use (model_stats/<model>.txt)\n"
           "  num_units        Number of parameter units\n"
           "  sharding_factor  Sharding factor for each unit (must divide the world size)\n"
           "  base_path        Base path for the repository (relative to the current directory)\n\n",
           prog);
    b200::print_options_help();
}

static void clear_timers() {
    timer_runtime.clear();
    timer_allgather.clear();
    timer_allgather_wait_fwd.clear();
    timer_allgather_wait_bwd.clear();
    timer_reduce_scatter.clear();
    timer_barrier.clear();
    exposed_wait_fwd_ms.clear();
    exposed_wait_bwd_ms.clear();
}

template <typename T> static int proxy_main(const b200::Cli &cli, const b200::World &w) {
    const std::string model_name = cli.positional[0];
    const unsigned num_units = (unsigned)atoi(cli.positional[1].c_str());
    const unsigned sharding_factor = (unsigned)atoi(cli.positional[2].c_str());
    if (num_units == 0 || sharding_factor == 0) {
        fprintf(stderr, "Error: num_units and sharding_factor must be positive\n");
        return 1;
    }
    if ((unsigned)w.size % sharding_factor != 0) {
        fprintf(stderr, "Error: world size %d is not a multiple of the sharding factor %u\n", w.size, sharding_factor);
        return 1;
    }
    const fs::path stats_file = get_dnnproxy_base_path(cli.positional[3].c_str()) / "model_stats" / (model_name + ".txt");
    if (!fs::exists(stats_file)) {
        if (w.rank == 0) std::cerr << "Error: model stats file does not exist: " << stats_file << "\n";
        return -1;
    }
    std::map<std::string, uint64_t> stats = get_model_stats(stats_file.string());
    const uint64_t total_model_size = stats["modelSize"];
    const unsigned local_batch_size = (unsigned)stats["batchSize"];
    const uint64_t fwd_rt_whole_model = stats["avgForwardTime"], bwd_rt_whole_model = stats["avgBackwardTime"];

    // unit u holds floor(M/U) (+1 for the first M mod U) parameters; its shard is the ceiling of 1/F of that
    std::vector<uint64_t> max_params_per_shard(num_units);
    uint64_t max_shard_size = 0;
    for (unsigned u = 0; u < num_units; u++) {
        const uint64_t unit = total_model_size / num_units + (u < total_model_size % num_units ? 1 : 0);
        max_params_per_shard[u] = unit / sharding_factor + (unit % sharding_factor ? 1 : 0);
        max_shard_size = std::max(max_shard_size, max_params_per_shard[u]);
    }
    if (max_shard_size * sharding_factor > 0x7fffffffull * (uint64_t)sharding_factor || max_shard_size > 0x7fffffffull) {
        fprintf(stderr, "Error: %llu elements per shard do not fit the interface's int count; use more units\n",
                (unsigned long long)max_shard_size);
        return 1;
    }

    const int my_device = set_local_device(w.local_rank, cli.devices.c_str());
    b200::check(dlnb_init(w.rank, w.size, my_device, w.session.c_str()), "dlnb_init");
    print_topology_graph();
    {
        char line[128];
        snprintf(line, sizeof(line), "Rank %d using device %d\n", w.rank, my_device);
        b200::print_all_ordered(line);
    }
    b200::apply_spin_cli(cli);

    // unit group: the F consecutive ranks sharing rank / F; replica group: ranks with equal rank % F
    const unsigned num_replicas = (unsigned)w.size / sharding_factor;
    std::vector<int> unit_group, replica_group;
    for (unsigned i = 0; i < sharding_factor; i++) unit_group.push_back((int)((unsigned)w.rank / sharding_factor * sharding_factor + i));
    for (unsigned i = 0; i < num_replicas; i++) replica_group.push_back((int)((unsigned)w.rank % sharding_factor + i * sharding_factor));
    ProxyBackend *unit_comm_proxy = new ProxyBackend(unit_group, (int)num_units, ProxyDtype<T>::value);
    ProxyBackend *allreduce_comm_proxy = new ProxyBackend(replica_group, (int)num_units, ProxyDtype<T>::value);
    unit_comm_proxy->set("algo", b200::algo_from_name(cli.algo));
    allreduce_comm_proxy->set("algo", b200::algo_from_name(cli.algo));
    if (cli.ctas > 0) {
        unit_comm_proxy->set("ctas", cli.ctas);
        allreduce_comm_proxy->set("ctas", cli.ctas);
    }

    Tensor<T, device> *allgather_buf = new Tensor<T, device>(max_shard_size * sharding_factor);
    std::vector<Tensor<T, device> *> shard_params(num_units), allreduce_params(num_units, nullptr);
    for (unsigned u = 0; u < num_units; u++) {
        shard_params[u] = new Tensor<T, device>(max_params_per_shard[u]);
        if (num_replicas > 1) allreduce_params[u] = new Tensor<T, device>(max_params_per_shard[u]);
        b200::check(dlnb_fill_pattern(shard_params[u]->data, max_params_per_shard[u], ProxyDtype<T>::value, DLNB_PAT_INT, 0, w.rank),
                    "fill");
    }
    const float fwd_rt_whole_unit = (float)fwd_rt_whole_model / num_units;
    const float bwd_rt_whole_unit = (float)bwd_rt_whole_model / num_units;
    double spin_overhead_us = 0, spin_slope = 1;
    b200::check(dlnb_spin_calibrate(&spin_overhead_us, &spin_slope), "spin calibration");
    b200::host_barrier();

    auto plain_iteration = [&]() {
        run_fsdp<T>(shard_params.data(), allgather_buf, allreduce_params.data(), fwd_rt_whole_unit, bwd_rt_whole_unit,
                    num_units, sharding_factor, max_params_per_shard.data(), num_replicas, unit_comm_proxy,
                    allreduce_comm_proxy);
    };
    in_graph_capture = cli.graph; // slot times are not host-visible inside a captured iteration
    auto iteration = b200::make_runner(plain_iteration, cli.graph);
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
    clear_timers();
#ifdef PROXY_LOOP
    b200::install_stop_handlers();
    const long bound = b200::loop_iters_from_env();
    for (long it = 0; (bound < 0 || it < bound) && !b200::stop_flag(); it++) {
        const bool keep = bound >= 0 || device_ms.size() < 4096;
        if (!keep) clear_timers();
        timer_runtime.start();
        b200::check(dlnb_iter_begin(), "iter begin");
        iteration();
        float ms = 0;
        b200::check(dlnb_iter_end(&ms), "iter end");
        timer_runtime.stop();
        if (keep) device_ms.push_back(ms);
    }
    if (bound < 0) {
        fprintf(stderr, "[rank %d] fsdp_loop stopped after %zu timed iterations, median %.3f ms (device)\n", w.rank,
                device_ms.size(), b200::median(device_ms));
        _exit(0);
    }
#else
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

    char host_name[256] = "localhost";
    gethostname(host_name, sizeof(host_name) - 1);
    const uint64_t allgather_msg_size = max_params_per_shard[0] * sharding_factor * sizeof(T);
    const uint64_t reducescatter_msg_size = max_params_per_shard[0] * sizeof(T);
    const uint64_t lower_bound_us = (uint64_t)(num_units - 1) * (uint64_t)fwd_rt_whole_unit + (uint64_t)num_units * (uint64_t)bwd_rt_whole_unit;
    const std::vector<float> device_ms_max = b200::max_over_ranks(device_ms);
    // device time of the most recent blocking collective = the last ReduceScatter (unit 0);
    // slots are not host-visible inside a graph replay
    const float last_rs_ms = cli.graph ? 0.f : unit_comm_proxy->slot_ms(-1);
    const int F = (int)sharding_factor;
    const double rs_bus = last_rs_ms <= 0 ? 0.0 : F > 1 ? (double)(F - 1) / F * (double)allgather_msg_size / (last_rs_ms * 1e-3) / 1e9
                                : 2.0 * (double)reducescatter_msg_size / (last_rs_ms * 1e-3) / 1e9;

    b200::Section fsdp("fsdp", "FSDP metrics");
    fsdp.per_rank.put("runtime", timer_runtime.vals)
        .put("allgather", timer_allgather.vals)
        .put("allgather_wait_fwd", timer_allgather_wait_fwd.vals)
        .put("allgather_wait_bwd", timer_allgather_wait_bwd.vals)
        .put("reduce_scatter", timer_reduce_scatter.vals)
        .put("allgather_exposed_wait_fwd_ms", exposed_wait_fwd_ms)
        .put("allgather_exposed_wait_bwd_ms", exposed_wait_bwd_ms)
        .put("barrier", timer_barrier.vals)
        .put("hostname", host_name)
        .put("rank", w.rank)
        .put("device_runtimes_ms", device_ms);
    fsdp.global.put("model_size_bytes", total_model_size * sizeof(T))
        .put("model_name", model_name)
        .put("world_size", w.size)
        .put("num_units", num_units)
        .put("sharding_factor", sharding_factor)
        .put("num_replicas", num_replicas)
        .put("local_batch_size", local_batch_size)
        .put("device", device == Device::CPU ? "CPU" : "GPU")
        .put("backend", unit_comm_proxy->get_name())
        .put("fwd_time_per_unit_us", fwd_rt_whole_unit)
        .put("bwd_time_per_unit_us", bwd_rt_whole_unit)
        .put("allgather_msg_size_bytes", allgather_msg_size)
        .put("reducescatter_msg_size_bytes", reducescatter_msg_size);
    if (num_replicas > 1) fsdp.global.put("allreduce_msg_size_bytes", max_params_per_shard[0] * sizeof(T));
    fsdp.global.put("dtype", ProxyDtype<T>::name)
        .put("arena", dlnb_arena_kind())
        .put("lower_bound_us", lower_bound_us)
        .put("device_iter_ms_max_over_ranks", device_ms_max)
        .put("device_iter_ms_median", b200::median(device_ms_max))
        .put("ratio_to_lower_bound", lower_bound_us ? b200::median(device_ms_max) * 1000.0 / (double)lower_bound_us : 0.0)
        .put("reduce_scatter_last_ms", last_rs_ms)
        .put("reduce_scatter_busbw_GBps", rs_bus)
        .put("allgather_first_ms", first_allgather_ms)
        .put("allgather_busbw_GBps", first_allgather_ms <= 0 ? 0.0 : F > 1 ? (double)(F - 1) / F * (double)allgather_msg_size / (first_allgather_ms * 1e-3) / 1e9
                                                            : 2.0 * (double)reducescatter_msg_size / (first_allgather_ms * 1e-3) / 1e9)
        .put("spin_overhead_us", spin_overhead_us)
        .put("spin_slope", spin_slope)
        .put("spin_ctas", (long)dlnb_get("spin_ctas"))
        .put("spin_threads", (long)dlnb_get("spin_threads"))
        .put("spin_mode", b200::spin_mode_name())
        .put("spin_coresident", (long)dlnb_get("spin_coresident"))
        .put("cuda_graph", cli.graph)
        // allgather_wait_* time Wait() from the moment the host reaches it. The reference's host sleeps through
        // the compute first (usleep), so its timers hold only the wait that is LEFT; here compute is asynchronous
        // and the timers also contain the overlapped part, unless DLNB_BLOCKING_COMPUTE=1 makes the host wait
        // for each spin like usleep does. The per-rank lists allgather_exposed_wait_{fwd,bwd}_ms hold the reference's meaning
        // either way: what was left of the allgather after the compute had ended, from device events.
        .put("wait_timers_include_overlap", !blocking_compute && !cli.graph)
        .put("kernel_launches", (uint64_t)dlnb_launch_count());
    fsdp.emit();

    unit_comm_proxy->finalize();
    allreduce_comm_proxy->finalize();
    b200::check(dlnb_finalize(), "dlnb_finalize");
    return 0;
}

int main(int argc, char *argv[]) {
    const b200::Cli cli = b200::parse_cli(argc, argv);
    const b200::World w = b200::world_from_env();
    if (argc < 5) {
        if (w.rank == 0)
            printf("Usage: dlnb_launch -n <world_size> ./fsdp <model_name> <num_units> <sharding_factor> <base_path>\n");
        return -1;
    }
    if (!cli.ok || cli.help || cli.positional.size() != 4) {
        print_help(argv[0]);
        return 1;
    }
    blocking_compute = getenv("DLNB_BLOCKING_COMPUTE") && atoi(getenv("DLNB_BLOCKING_COMPUTE")) != 0;
    switch (b200::dtype_from_name(cli.dtype)) {
    case DLNB_F32: return proxy_main<float>(cli, w);
    case DLNB_F16: return proxy_main<proxy_half>(cli, w);
    case DLNB_BF16: return proxy_main<proxy_bfloat16>(cli, w);
    }
    fprintf(stderr, "Error: unknown dtype '%s' (fp32 | fp16 | bf16)\n", cli.dtype.c_str());
    return 1;
}
