// dlnb_launch — starts the `n` single-GPU processes of one box (stand-in for mpirun/srun,
// which the target image does not have).
//
//   dlnb_launch -n <ranks> <program> [args...]
//
// Every child gets DLNB_RANK, DLNB_WORLD, DLNB_LOCAL_RANK and a fresh DLNB_SESSION (the name
// of the rendezvous the C ABI's dlnb_init joins). SIGINT/SIGTERM are forwarded; if a rank
// fails the others are terminated and its exit code is returned.
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/wait.h>
#include <unistd.h>

#include <string>
#include <vector>

static std::vector<pid_t> kids;
static void forward(int sig) {
    for (pid_t k : kids)
        if (k > 0) kill(k, sig);
}

int main(int argc, char **argv) {
    int n = 1, i = 1;
    for (; i < argc && argv[i][0] == '-'; i += 2) {
        if ((!strcmp(argv[i], "-n") || !strcmp(argv[i], "-np")) && i + 1 < argc) n = atoi(argv[i + 1]);
        else { i = argc; break; }
    }
    if (i >= argc || n < 1 || n > 8) {
        fprintf(stderr, "usage: dlnb_launch -n <ranks 1..8> <program> [args...]\n");
        return 2;
    }
    const std::string session = "launch-" + std::to_string(getpid());
    struct sigaction sa;
    memset(&sa, 0, sizeof(sa));
    sa.sa_handler = forward;
    sigaction(SIGINT, &sa, nullptr);
    sigaction(SIGTERM, &sa, nullptr);
    for (int r = 0; r < n; ++r) {
        pid_t p = fork();
        if (p < 0) { perror("fork"); forward(SIGTERM); return 2; }
        if (p == 0) {
            setenv("DLNB_RANK", std::to_string(r).c_str(), 1);
            setenv("DLNB_LOCAL_RANK", std::to_string(r).c_str(), 1);
            setenv("DLNB_WORLD", std::to_string(n).c_str(), 1);
            setenv("DLNB_SESSION", session.c_str(), 1);
            execvp(argv[i], argv + i);
            perror(argv[i]);
            _exit(127);
        }
        kids.push_back(p);
    }
    int rc = 0, left = n;
    while (left > 0) {
        int st = 0;
        pid_t p = wait(&st);
        if (p < 0) { if (errno == EINTR) continue; break; }
        --left;
        for (pid_t &k : kids) if (k == p) k = -1;
        const int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st);
        if (code != 0 && rc == 0) {
            rc = code;
            if (WIFSIGNALED(st) || code == 1) forward(SIGTERM); // a rank died: do not leave the others waiting
        }
    }
    return rc;
}
