// CPU-only check of the rendezvous used for VMM handle exchange: N forked processes do an
// allgather, a barrier and an fd exchange (each contributes a pipe whose write end it keeps;
// after the exchange every rank reads every other rank's message through the received fds).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/wait.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "../../dlnetbench_b200/csrc/bootstrap.hpp"

static int rank_main(int rank, int world, const std::string &session) {
    dlnb::Bootstrap b;
    std::string e = b.connect(rank, world, session, 20);
    if (!e.empty()) { fprintf(stderr, "connect: %s\n", e.c_str()); return 1; }
    std::vector<int> all(world);
    int mine = rank * 7 + 1;
    if (!(e = b.allgather(&mine, sizeof(int), all.data())).empty()) return 2;
    for (int r = 0; r < world; ++r) if (all[r] != r * 7 + 1) return 3;
    int pfd[2];
    if (pipe(pfd)) return 4;
    std::vector<int> fds;
    if (!(e = b.exchange_fds(pfd[0], fds)).empty()) { fprintf(stderr, "fds: %s\n", e.c_str()); return 5; }
    char msg[32];
    snprintf(msg, sizeof(msg), "hello from %d", rank);
    // one copy per reader (everyone reads from everyone's pipe once)
    for (int k = 0; k < world; ++k) if (write(pfd[1], msg, 32) != 32) return 6;
    if (!(e = b.barrier()).empty()) return 7;
    for (int r = 0; r < world; ++r) {
        char got[32] = {0}, want[32];
        if (read(fds[r], got, 32) != 32) return 8;
        snprintf(want, sizeof(want), "hello from %d", r);
        if (strcmp(got, want)) { fprintf(stderr, "rank %d: fd %d gave '%s'\n", rank, r, got); return 9; }
    }
    if (!(e = b.barrier()).empty()) return 10;
    return 0;
}

int main(int argc, char **argv) {
    const int world = argc > 1 ? atoi(argv[1]) : 3;
    const std::string session = "bt-" + std::to_string(getpid());
    std::vector<pid_t> kids;
    for (int r = 0; r < world; ++r) {
        pid_t p = fork();
        if (p == 0) _exit(rank_main(r, world, session));
        kids.push_back(p);
    }
    int rc = 0;
    for (pid_t p : kids) {
        int st = 0;
        waitpid(p, &st, 0);
        if (!WIFEXITED(st) || WEXITSTATUS(st)) rc = WIFEXITED(st) ? WEXITSTATUS(st) : 99;
    }
    printf(rc ? "FAIL %d\n" : "OK\n", rc);
    return rc;
}
