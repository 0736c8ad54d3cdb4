// cpp/netcommunicators.hpp — topology printer, and the communicator interface re-exported.
//
// In the reference this header holds only the SLURM topology ASCII printer
// (allgather_strings :19, print_topology_graph :142); the communicator interface lives in
// proxy_classes.hpp. Both are reachable from here so that "the communicator interface in
// netcommunicators.hpp" holds literally for the b200 build.
#ifndef NETCOMMUNICATORS_HPP
#define NETCOMMUNICATORS_HPP

#include <dlnb.h>

#include <cstdlib>
#include <map>
#include <string>
#include <vector>

#include "b200/host_runtime.hpp"
#include "proxy_classes.hpp"

// Every rank contributes one string; all ranks get the list in rank order
// (the reference does MPI_Allgather of lengths + MPI_Allgatherv, netcommunicators.hpp:19-77).
inline std::vector<std::string> allgather_strings(const std::string &mine) {
    return b200::host_allgather_strings(mine);
}

// Rank 0 draws where the ranks sit. Uses SLURM_TOPOLOGY_ADDR ("switch. ... .node") when set;
// otherwise the ranks of this launch share one NVSwitch box, which is what gets drawn.
inline void print_topology_graph() {
    const char *env = std::getenv("SLURM_TOPOLOGY_ADDR");
    char host[256] = "localhost";
    gethostname(host, sizeof(host) - 1);
    const std::string addr = env ? std::string(env) : std::string("NVSwitch.") + host;
    const std::vector<std::string> all = allgather_strings(addr + "#" + std::to_string(dlnb_device()));
    if (dlnb_rank() != 0) return;
    // group ranks by everything before the '#'
    std::map<std::string, std::vector<std::string>> by_node;
    for (size_t r = 0; r < all.size(); ++r) {
        const size_t h = all[r].rfind('#');
        by_node[all[r].substr(0, h)].push_back("r" + std::to_string(r) + ":gpu" + all[r].substr(h + 1));
    }
    std::string out = "=== Topology Graph ===\n";
    for (const auto &kv : by_node) {
        std::string path = kv.first, line;
        size_t depth = 0, pos = 0;
        while (true) {
            const size_t dot = path.find('.', pos);
            const std::string level = path.substr(pos, dot == std::string::npos ? std::string::npos : dot - pos);
            out += std::string(depth * 2, ' ') + "+-[" + level + "]\n";
            ++depth;
            if (dot == std::string::npos) break;
            pos = dot + 1;
        }
        for (const std::string &g : kv.second) line += " " + g;
        out += std::string(depth * 2, ' ') + "|" + line + "   (NVLink 5, all-to-all through NVSwitch)\n";
    }
    out += "======================\n";
    b200::print_once(out);
}

#endif // NETCOMMUNICATORS_HPP
