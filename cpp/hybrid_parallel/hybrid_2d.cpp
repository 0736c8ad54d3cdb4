// hybrid_2d / hybrid_2d_loop — data + pipeline parallel (GPipe) training proxy, b200 build.
//
// Same CLI, inputs and output section as the reference's cpp/hybrid_parallel/hybrid_2d.cpp
// (`hybrid_2d <model> <num_stages> <num_microbatches> <base_path> [-w][-r][-d][-m][-h]`, section
// "dp_pp" with the keys of hybrid_2d.cpp:443-463). World rank r sits in pipeline r / S at stage
// r % S; the S ranks of a pipeline exchange activations (forward) and gradients (backward) per
// micro-batch with blocking send / recv, then the ranks holding the same stage allreduce that
// stage's gradients (hybrid_2d.cpp:90-169). Here send / recv are rendezvous puts over peer HBM
// and the per-micro-batch compute is the %globaltimer spin kernel.
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <vector>

#include "../data_types.hpp"
#include "../netcommunicators.hpp"
#include "../proxy_classes.hpp"
#include "../utils.hpp"

constexpr Device device = Device::GPU;

static b200::HostTimer timer_runtime, timer_pp_comm, timer_dp_comm;

static inline void emulate_compute(uint64_t usec) { b200::check(dlnb_compute_spin((double)usec), "compute spin"); }

// One GPipe iteration: all micro-batches forward, then all backward, then the DP allreduce.
template <typename T>
int run_data_pipe_parallel(int num_microbatches, int stage_id, int num_stage, uint64_t pipe_msg_size, uint64_t fwd_rt,
                           uint64_t bwd_rt, Tensor<T, device> *grad_ptr, Tensor<T, device> *sum_grad_ptr,
                           uint64_t dp_allreduce_size, Tensor<T, device> *fwd_send_buff, Tensor<T, device> *fwd_recv_buff,
                           Tensor<T, device> *bwd_send_buff, Tensor<T, device> *bwd_recv_buff,
                           ProxyCommunicator *dp_communicator, ProxyCommunicator *pp_communicator) {
    const bool first = stage_id == 0, last = stage_id == num_stage - 1;
    auto timed = [&](auto &&call) {
        timer_pp_comm.start();
        call();
        timer_pp_comm.stop();
    };
    for (int mb = 0; mb < num_microbatches; mb++) { // forward: activations flow down the pipeline
        if (!first) timed([&] { pp_communicator->recv(fwd_recv_buff->data, (int)pipe_msg_size, stage_id - 1); });
        emulate_compute(fwd_rt);
        if (!last) timed([&] { pp_communicator->send(fwd_send_buff->data, (int)pipe_msg_size, stage_id + 1); });
    }
    for (int mb = 0; mb < num_microbatches; mb++) { // backward: gradients flow back up
        if (!last) timed([&] { pp_communicator->recv(bwd_recv_buff->data, (int)pipe_msg_size, stage_id + 1); });
        emulate_compute(bwd_rt);
        if (!first) timed([&] { pp_communicator->send(bwd_send_buff->data, (int)pipe_msg_size, stage_id - 1); });
    }
    timer_dp_comm.start();
    proxy_allreduce(dp_communicator, grad_ptr->data, sum_grad_ptr->data, dp_allreduce_size); // may exceed the interface's int
    timer_dp_comm.stop();
    return 0;
}

static void print_help(const char *prog) {
    printf("Usage: %s <model> <num_stages> <num_microbatches> <base_path> [options]\n\n"
           "Arguments:\n"
           "  model             Name of the model to use (model_stats/<model>.txt, models/<family>.json)\n"
           "  num_stages        Number of pipeline stages\n"
           "  num_microbatches  Number of microbatches\n"
           "  base_path         Base path for the repository (relative to the current directory)\n\n",
           prog);
    b200::print_options_help();
}

template <typename T> static int proxy_main(const b200::Cli &cli, const b200::World &w) {
    const std::string model_name = cli.positional[0];
    const int num_stage = atoi(cli.positional[1].c_str()), num_microbatches = atoi(cli.positional[2].c_str());
    if (num_stage <= 0 || num_microbatches <= 0) {
        fprintf(stderr, "Error: num_stages and num_microbatches must be positive\n");
        return 1;
    }
    const fs::path repo = get_dnnproxy_base_path(cli.positional[3].c_str());
    const fs::path stats_file = repo / "model_stats" / (model_name + ".txt");
    // models/<family>.json: the stats name without its two trailing _<batch>_<dtype> fields
    std::string family = model_name.substr(0, model_name.find_last_of('_'));
    family = family.substr(0, family.find_last_of('_'));
    if (!fs::exists(stats_file)) {
        std::cerr << "Error: model stats file does not exist: " << stats_file << "\n";
        return -1;
    }
    const unsigned num_layers = count_layers((repo / "models" / (family + ".json")).string());
    std::map<std::string, uint64_t> stats = get_model_stats(stats_file.string());
    const unsigned local_batch_size = (unsigned)stats["batchSize"];
    const uint64_t total_model_size = stats["modelSize"];
    // per stage, then per micro-batch (integer divisions, as the reference)
    const uint64_t fwd_rt_per_microbatch = stats["avgForwardTime"] / num_stage / num_microbatches;
    const uint64_t bwd_rt_per_microbatch = stats["avgBackwardTime"] / num_stage / num_microbatches;
    const uint64_t samples_per_microbatch = local_batch_size / num_microbatches;
    const uint64_t pipe_msg_size = (uint64_t)((unsigned)stats["sequenceLength"] * (unsigned)stats["embeddedDim"] * samples_per_microbatch);
    const uint64_t dp_allreduce_size = total_model_size / num_stage;
    if (num_layers % num_stage != 0 || local_batch_size % num_microbatches != 0 || w.size % num_stage != 0) {
        fprintf(stderr, "Error: layers (%u) and world size (%d) must be multiples of the stages (%d), the batch (%u) of the "
                        "microbatches (%d)\n", num_layers, w.size, num_stage, local_batch_size, num_microbatches);
        return 1;
    }
    // (the per-stage gradient allreduce may exceed int: it goes through proxy_allreduce's 64-bit entry)
    if (pipe_msg_size > 0x7fffffffull) {
        fprintf(stderr, "Error: stage-to-stage message of %llu elements does not fit the interface's int count\n",
                (unsigned long long)pipe_msg_size);
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

    // DP group: the ranks with my stage (rank % S); PP group: my pipeline (rank / S), in stage order
    const int dp_size = w.size / num_stage, stage_id = w.rank % num_stage;
    std::vector<int> dp_group, pp_group;
    for (int i = 0; i < dp_size; i++) dp_group.push_back(stage_id + i * num_stage);
    for (int i = 0; i < num_stage; i++) pp_group.push_back(w.rank / num_stage * num_stage + i);
    ProxyBackend *dp_communicator = new ProxyBackend(dp_group, 1, ProxyDtype<T>::value);
    ProxyBackend *pp_communicator = new ProxyBackend(pp_group, 1, ProxyDtype<T>::value);
    dp_communicator->set("algo", b200::algo_from_name(cli.algo));
    if (cli.ctas > 0) {
        dp_communicator->set("ctas", cli.ctas);
        pp_communicator->set("ctas", cli.ctas);
    }

    Tensor<T, device> *grad_ptr = new Tensor<T, device>(dp_allreduce_size), *sum_grad_ptr = new Tensor<T, device>(dp_allreduce_size);
    Tensor<T, device> *fwd_send_buff = new Tensor<T, device>(pipe_msg_size), *fwd_recv_buff = new Tensor<T, device>(pipe_msg_size);
    Tensor<T, device> *bwd_send_buff = new Tensor<T, device>(pipe_msg_size), *bwd_recv_buff = new Tensor<T, device>(pipe_msg_size);
    b200::check(dlnb_fill_pattern(grad_ptr->data, dp_allreduce_size, ProxyDtype<T>::value, DLNB_PAT_INT, 0, w.rank), "fill");
    b200::check(dlnb_fill_pattern(fwd_send_buff->data, pipe_msg_size, ProxyDtype<T>::value, DLNB_PAT_INT, 1, w.rank), "fill");
    b200::check(dlnb_fill_pattern(bwd_send_buff->data, pipe_msg_size, ProxyDtype<T>::value, DLNB_PAT_INT, 2, w.rank), "fill");
    double spin_overhead_us = 0, spin_slope = 1;
    b200::check(dlnb_spin_calibrate(&spin_overhead_us, &spin_slope), "spin calibration");
    b200::host_barrier();

    auto iteration = [&]() {
        run_data_pipe_parallel<T>(num_microbatches, stage_id, num_stage, pipe_msg_size, fwd_rt_per_microbatch,
                                  bwd_rt_per_microbatch, grad_ptr, sum_grad_ptr, dp_allreduce_size, fwd_send_buff, fwd_recv_buff,
                                  bwd_send_buff, bwd_recv_buff, dp_communicator, pp_communicator);
    };
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
    timer_pp_comm.clear();
    timer_dp_comm.clear();
#ifdef PROXY_LOOP
    b200::install_stop_handlers();
    const long bound = b200::loop_iters_from_env();
    for (long it = 0; (bound < 0 || it < bound) && !b200::stop_flag(); it++) {
        if (bound < 0 && device_ms.size() >= 4096) { timer_pp_comm.clear(); timer_dp_comm.clear(); timer_runtime.clear(); device_ms.clear(); }
        timer_runtime.start();
        b200::check(dlnb_iter_begin(), "iter begin");
        iteration();
        float ms = 0;
        b200::check(dlnb_iter_end(&ms), "iter end");
        timer_runtime.stop();
        device_ms.push_back(ms);
    }
    if (bound < 0) {
        fprintf(stderr, "[rank %d] hybrid_2d_loop stopped after %zu timed iterations, median %.3f ms (device)\n", w.rank,
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

    // middle stages time a recv and a send per micro-batch and direction: report their sum, one
    // entry per micro-batch and direction, like the reference does (hybrid_2d.cpp:414-436)
    std::vector<float> pp_times = timer_pp_comm.vals;
    if (stage_id > 0 && stage_id < num_stage - 1) {
        std::vector<float> merged;
        for (size_t i = 0; i + 1 < pp_times.size(); i += 2) merged.push_back(pp_times[i] + pp_times[i + 1]);
        pp_times.swap(merged);
    }
    // GPipe critical path of pure emulated compute: (M + S - 1) slots forward and backward
    const uint64_t lower_bound_us = (uint64_t)(num_microbatches + num_stage - 1) * (fwd_rt_per_microbatch + bwd_rt_per_microbatch);
    const std::vector<float> device_ms_max = b200::max_over_ranks(device_ms);
    const float dp_ms = dp_communicator->slot_ms(-1);
    const double wire = dp_size > 1 ? 2.0 * (dp_size - 1) / dp_size * (double)dp_allreduce_size * sizeof(T)
                                    : 2.0 * (double)dp_allreduce_size * sizeof(T);
    char host_name[256] = "localhost";
    gethostname(host_name, sizeof(host_name) - 1);

    b200::Section out("dp_pp", "Data + Pipeline Parallelism");
    out.global.put("model_name", model_name)
        .put("num_stages", num_stage)
        .put("num_microbatches", num_microbatches)
        .put("samples_per_microbatch", samples_per_microbatch)
        .put("local_batch_size", local_batch_size)
        .put("global_batch_size", (unsigned)w.size * local_batch_size / (unsigned)num_stage)
        .put("world_size", w.size)
        .put("dp_size", dp_size)
        .put("fwd_rt_per_microbatch", fwd_rt_per_microbatch)
        .put("bwd_rt_per_microbatch", bwd_rt_per_microbatch)
        .put("total_model_size_params", total_model_size)
        .put("pipe_msg_size_bytes", pipe_msg_size * sizeof(T))
        .put("dp_allreduce_size_bytes", dp_allreduce_size * sizeof(T))
        .put("device", device == Device::CPU ? "CPU" : "GPU")
        .put("backend", dp_communicator->get_name())
        .put("dtype", ProxyDtype<T>::name)
        .put("arena", dlnb_arena_kind())
        .put("lower_bound_us", lower_bound_us)
        .put("device_iter_ms_max_over_ranks", device_ms_max)
        .put("device_iter_ms_median", b200::median(device_ms_max))
        .put("ratio_to_lower_bound", lower_bound_us ? b200::median(device_ms_max) * 1000.0 / (double)lower_bound_us : 0.0)
        .put("dp_allreduce_last_ms", dp_ms)
        .put("dp_allreduce_busbw_GBps", dp_ms > 0 ? wire / (dp_ms * 1e-3) / 1e9 : 0.0)
        .put("kernel_launches", (uint64_t)dlnb_launch_count());
    out.per_rank.put("runtimes", timer_runtime.vals)
        .put("pp_comm_time", pp_times)
        .put("dp_comm_time", timer_dp_comm.vals)
        .put("hostname", host_name)
        .put("stage_id", stage_id)
        .put("device_runtimes_ms", device_ms);
    out.emit();

    dp_communicator->finalize();
    pp_communicator->finalize();
    b200::check(dlnb_finalize(), "dlnb_finalize");
    return 0;
}

int main(int argc, char *argv[]) {
    const b200::Cli cli = b200::parse_cli(argc, argv);
    if (!cli.ok || cli.help || cli.positional.size() != 4) {
        print_help(argv[0]);
        return 1;
    }
    const b200::World w = b200::world_from_env();
    switch (b200::dtype_from_name(cli.dtype)) {
    case DLNB_F32: return proxy_main<float>(cli, w);
    case DLNB_F16: return proxy_main<proxy_half>(cli, w);
    case DLNB_BF16: return proxy_main<proxy_bfloat16>(cli, w);
    }
    fprintf(stderr, "Error: unknown dtype '%s' (fp32 | fp16 | bf16)\n", cli.dtype.c_str());
    return 1;
}
