/* mpiexec for mpishim: `mpiexec -n <P> [-f machines] prog args...` forks P ranks on this node
 * that share one POSIX shared-memory file.  HARNESS ONLY (see mpi.h). */
#define _GNU_SOURCE
#include <fcntl.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#define RING_BYTES (1u << 20)
#define SHM_HEADER 4096 /* >= sizeof(Shm) header incl. the reduce row */

int main(int argc, char **argv)
{
    int n = 1, i = 1;
    while (i < argc && argv[i][0] == '-') {
        if ((!strcmp(argv[i], "-n") || !strcmp(argv[i], "-np")) && i + 1 < argc) { n = atoi(argv[i + 1]); i += 2; }
        else if (!strcmp(argv[i], "-f") && i + 1 < argc) i += 2; /* hostfile: single node, ignored */
        else ++i;
    }
    if (i >= argc || n < 1 || n > 64) {
        fprintf(stderr, "usage: mpiexec -n <ranks 1..64> prog args...\n");
        return 2;
    }
    char name[64];
    snprintf(name, sizeof name, "/mpishim_%d", (int)getpid());
    int fd = shm_open(name, O_CREAT | O_RDWR | O_EXCL, 0600);
    if (fd < 0) { perror("shm_open"); return 1; }
    size_t bytes = SHM_HEADER + ((size_t)RING_BYTES + 128) * (size_t)n * (size_t)n;
    if (ftruncate(fd, (off_t)bytes) != 0) { perror("ftruncate"); shm_unlink(name); return 1; }
    close(fd);
    char sz[16];
    snprintf(sz, sizeof sz, "%d", n);
    setenv("MPISHIM_SHM", name, 1);
    setenv("MPISHIM_SIZE", sz, 1);
    pid_t kids[64];
    for (int r = 0; r < n; ++r) {
        pid_t pid = fork();
        if (pid < 0) { perror("fork"); return 1; }
        if (pid == 0) {
            char rk[16];
            snprintf(rk, sizeof rk, "%d", r);
            setenv("MPISHIM_RANK", rk, 1);
            execvp(argv[i], argv + i);
            perror("execvp");
            _exit(127);
        }
        kids[r] = pid;
    }
    int status = 0, left = n, failed = 0;
    while (left > 0) {
        int st;
        pid_t p = wait(&st);
        if (p < 0) break;
        --left;
        if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) {
            if (!failed) /* one rank died: MPI_ERRORS_ARE_FATAL - take the job down */
                for (int r = 0; r < n; ++r)
                    if (kids[r] != p) kill(kids[r], SIGTERM);
            failed = 1;
            status = WIFEXITED(st) ? WEXITSTATUS(st) : 1;
        }
    }
    shm_unlink(name);
    return failed ? (status ? status : 1) : 0;
}
