// oracle/compat/mpirun.cpp — TEST INFRASTRUCTURE (oracle). Stand-in for `mpirun -n K prog args`
// for programs built against oracle/compat/mpi.h: creates the shared segment, forks K ranks
// with DLNB_MPI_RANK / DLNB_MPI_SIZE / DLNB_MPI_SHM set, waits, propagates failure.
#include <fcntl.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "mpi.h"

int main(int argc, char **argv) {
    int n = 1, i = 1;
    while (i < argc && argv[i][0] == '-') {
        if ((!strcmp(argv[i], "-n") || !strcmp(argv[i], "-np")) && i + 1 < argc) { n = atoi(argv[i + 1]); i += 2; }
        else { fprintf(stderr, "usage: dlnb_mpirun -n <ranks> <program> [args...]\n"); return 2; }
    }
    if (i >= argc || n < 1 || n > dlnb_mpi::MAX_RANKS) { fprintf(stderr, "usage: dlnb_mpirun -n <ranks> <program> [args...]\n"); return 2; }
    std::string name = "/dlnb_mpi_" + std::to_string(getpid());
    size_t bytes = dlnb_mpi::shared_bytes(n);
    int fd = shm_open(name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) { perror("shm_open"); return 2; }
    if (ftruncate(fd, (off_t)bytes) != 0) { perror("ftruncate"); shm_unlink(name.c_str()); return 2; }
    close(fd); /* zero-filled by the kernel: counters start at 0 */
    std::vector<pid_t> kids;
    for (int r = 0; r < n; ++r) {
        pid_t p = fork();
        if (p == 0) {
            setenv("DLNB_MPI_RANK", std::to_string(r).c_str(), 1);
            setenv("DLNB_MPI_SIZE", std::to_string(n).c_str(), 1);
            setenv("DLNB_MPI_SHM", name.c_str(), 1);
            execvp(argv[i], argv + i);
            perror("execvp");
            _exit(127);
        }
        kids.push_back(p);
    }
    int rc = 0, left = n;
    while (left > 0) {
        int st = 0;
        pid_t p = wait(&st);
        if (p < 0) break;
        left--;
        int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st);
        /* the reference's mains return -1 / 1 on argument errors on every rank: not a crash */
        if (code != 0 && rc == 0) rc = code;
        if (WIFSIGNALED(st))
            for (pid_t k : kids) if (k != p) kill(k, SIGTERM);
    }
    shm_unlink(name.c_str());
    return rc;
}
