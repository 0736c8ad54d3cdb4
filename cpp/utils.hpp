// cpp/utils.hpp — paths, device selection, model-stats parsing, run estimation, message stats.
//
// Same entry points as the reference's cpp/utils.hpp (get_dnnproxy_base_path :44, parse_devices
// :62, set_local_device :74, estimate_runs :121, compute_msg_stats :145, extract_value :167,
// get_model_stats :200), written for a build without MPI or nlohmann: cross-rank steps use the
// C ABI's rendezvous. `count_layers` (:279) serves the hybrid proxies.
#ifndef UTILS_HPP
#define UTILS_HPP

#include <dlnb.h>
#include <stdint.h>

#include <cmath>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace fs = std::filesystem;

// `input_path` is relative to the current directory, as in the reference (utils.hpp:53).
inline fs::path get_dnnproxy_base_path(const char *input_path) {
    const fs::path base = fs::current_path() / (input_path ? input_path : "");
    if (!fs::is_directory(base)) {
        std::cerr << "Error: DNNProxy folder does not exist at: " << base << "\n";
        throw std::runtime_error("DNNProxy folder not found");
    }
    return base;
}

inline std::vector<int> parse_devices(const std::string &s) {
    std::vector<int> ids;
    std::stringstream ss(s);
    for (std::string tok; std::getline(ss, tok, ',');)
        if (!tok.empty()) ids.push_back(std::stoi(tok));
    return ids;
}

// Device index for this process: entry `local_rank` of the -d list, or local_rank itself.
// (The reference derives local_rank from MPI_Comm_split_type; one box => local == launcher rank.)
inline int set_local_device(int local_rank, const char *devices_str) {
    const std::vector<int> list = parse_devices(devices_str ? devices_str : "");
    if (list.empty()) return local_rank;
    if (local_rank >= (int)list.size())
        throw std::runtime_error("device list shorter than the number of local ranks");
    return list[(size_t)local_rank];
}

// Runs needed to fill `min_exectime` seconds, from the warm-up times (first two discarded).
// The reference sums the per-rank means and divides by the warm-up COUNT (utils.hpp:127-128);
// here it is the mean over ranks, and fewer than 3 warm-ups fall back to all of them.
inline int estimate_runs(const std::vector<float> &warmup_times, uint64_t min_exectime) {
    const size_t skip = warmup_times.size() > 2 ? 2 : 0;
    float mine = 0.f;
    for (size_t i = skip; i < warmup_times.size(); ++i) mine += warmup_times[i];
    mine = warmup_times.size() > skip ? mine / (float)(warmup_times.size() - skip) : 0.f;
    std::vector<float> all((size_t)dlnb_world());
    if (dlnb_host_allgather(&mine, sizeof(float), all.data()) != 0) throw std::runtime_error(dlnb_last_error());
    float mean = 0.f;
    for (float t : all) mean += t;
    mean /= (float)all.size();
    if (!(mean > 0.f)) return 1;
    return std::max(1, (int)std::ceil((float)min_exectime / mean));
}

// Mean and population standard deviation of message sizes (elements), in float like the reference.
inline std::pair<float, float> compute_msg_stats(const std::vector<uint64_t> &sizes, unsigned sharding_multiplier = 1) {
    float avg = 0.0f;
    for (uint64_t s : sizes) avg += s * sharding_multiplier;
    avg /= sizes.size();
    float var = 0.0f;
    for (uint64_t s : sizes) {
        const float d = s * sharding_multiplier - avg;
        var += d * d;
    }
    return {avg, std::sqrt(var / sizes.size())};
}

// Text after the first ':' of a "key:value" line, trimmed; "" when there is no ':'.
inline std::string extract_value(const std::string &line) {
    const size_t colon = line.find(':');
    if (colon == std::string::npos) return "";
    const char *ws = " \t\r\n";
    const size_t b = line.find_first_not_of(ws, colon + 1);
    if (b == std::string::npos) return "";
    const size_t e = line.find_last_not_of(ws);
    return line.substr(b, e - b + 1);
}

// model_stats/<name>.txt is read POSITIONALLY: 12 lines, the two averages are
// floating-point text truncated to integers, everything else an unsigned integer prefix.
inline std::map<std::string, uint64_t> get_model_stats(const std::string &filename) {
    std::ifstream in(filename);
    if (!in.is_open()) throw std::runtime_error("Could not open model stats file: " + filename);
    static const char *keys[12] = {"forwardFlops", "backwardFlops", "modelSize", "nonExpertModelSize",
                                   "avgForwardTime", "avgBackwardTime", "batchSize", "ffn_avgForwardTime",
                                   "ffn_avgBackwardTime", "experts", "sequenceLength", "embeddedDim"};
    std::map<std::string, uint64_t> stats;
    std::string line;
    for (int k = 0; k < 12; ++k) {
        if (!std::getline(in, line)) throw std::runtime_error("model stats file too short: " + filename);
        const std::string v = extract_value(line);
        stats[keys[k]] = (k == 4 || k == 5) ? (uint64_t)std::stod(v) : std::stoull(v);
    }
    return stats;
}

// Number of layers of a model: num_encoder_blocks + num_decoder_blocks of models/<name>.json
// (reference: count_layers, utils.hpp:279-294, there through nlohmann::json). The files are flat
// objects of integers, so the two keys are read directly.
inline unsigned count_layers(const std::string &filename) {
    std::ifstream in(filename);
    if (!in.is_open()) throw std::runtime_error("Could not open model architecture file: " + filename);
    std::stringstream ss;
    ss << in.rdbuf();
    const std::string text = ss.str();
    auto value_of = [&](const std::string &key) -> unsigned {
        const size_t k = text.find("\"" + key + "\"");
        if (k == std::string::npos) return 0;
        const size_t colon = text.find(':', k);
        if (colon == std::string::npos) return 0;
        return (unsigned)std::strtoul(text.c_str() + colon + 1, nullptr, 10);
    };
    return value_of("num_encoder_blocks") + value_of("num_decoder_blocks");
}

#endif // UTILS_HPP
