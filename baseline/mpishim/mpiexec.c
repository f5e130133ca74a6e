/*
 * mpiexec for the single-node MPI shim:  mpiexec -n P program [args...]
 * Creates one Unix stream socket pair per pair of ranks, forks P processes and
 * tells each one its rank, the world size and the descriptor that leads to
 * every peer (environment MPISHIM_RANK / MPISHIM_SIZE / MPISHIM_FDS).  If a rank
 * exits with an error (MPI_Abort) the others are terminated.
 */
#define _GNU_SOURCE
#include <errno.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/resource.h>
#include <sys/socket.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

int main(int argc, char** argv) {
  int n = 1, a = 1;
  while (a < argc && argv[a][0] == '-') {
    if ((!strcmp(argv[a], "-n") || !strcmp(argv[a], "-np")) && a + 1 < argc) { n = atoi(argv[a + 1]); a += 2; }
    else { fprintf(stderr, "mpiexec (shim): unknown option %s\n", argv[a]); return 2; }
  }
  if (a >= argc || n < 1 || n > 64) { fprintf(stderr, "usage: mpiexec -n P program [args]   (1 <= P <= 64)\n"); return 2; }
  struct rlimit rl;
  if (getrlimit(RLIMIT_NOFILE, &rl) == 0 && rl.rlim_cur < rl.rlim_max) { rl.rlim_cur = rl.rlim_max; setrlimit(RLIMIT_NOFILE, &rl); }

  /* fd[i][j]: descriptor rank i uses to talk to rank j */
  int (*fd)[64] = calloc((size_t)n, sizeof *fd);
  for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) fd[i][j] = -1;
  for (int i = 0; i < n; ++i)
    for (int j = i + 1; j < n; ++j) {
      int sv[2];
      if (socketpair(AF_UNIX, SOCK_STREAM, 0, sv) != 0) { perror("socketpair"); return 1; }
      int sz = 4 << 20;  /* roomy buffers: eager sends of whole strips complete without the receiver */
      setsockopt(sv[0], SOL_SOCKET, SO_SNDBUF, &sz, sizeof sz); setsockopt(sv[1], SOL_SOCKET, SO_SNDBUF, &sz, sizeof sz);
      setsockopt(sv[0], SOL_SOCKET, SO_RCVBUF, &sz, sizeof sz); setsockopt(sv[1], SOL_SOCKET, SO_RCVBUF, &sz, sizeof sz);
      fd[i][j] = sv[0]; fd[j][i] = sv[1];
    }
  pid_t* pid = calloc((size_t)n, sizeof *pid);
  for (int r = 0; r < n; ++r) {
    pid[r] = fork();
    if (pid[r] < 0) { perror("fork"); return 1; }
    if (pid[r] == 0) {
      for (int i = 0; i < n; ++i) if (i != r) for (int j = 0; j < n; ++j) if (fd[i][j] >= 0) close(fd[i][j]);
      char buf[32], list[64 * 12] = "";
      snprintf(buf, sizeof buf, "%d", r); setenv("MPISHIM_RANK", buf, 1);
      snprintf(buf, sizeof buf, "%d", n); setenv("MPISHIM_SIZE", buf, 1);
      for (int j = 0; j < n; ++j) { snprintf(buf, sizeof buf, "%s%d", j ? "," : "", fd[r][j]); strcat(list, buf); }
      setenv("MPISHIM_FDS", list, 1);
      execvp(argv[a], &argv[a]);
      perror(argv[a]);
      _exit(127);
    }
  }
  for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) if (fd[i][j] >= 0) close(fd[i][j]);
  /* The first failing rank decides the exit code.  The others get a short grace
   * period to finish what they were printing (a program may report its error on
   * one rank while another trips over the same misconfiguration a moment
   * earlier), then SIGTERM. */
  int rc = 0, left = n, killed = 0;
  struct timespec t_fail = {0, 0};
  char* done = calloc((size_t)n, 1);
  while (left > 0) {
    int st; pid_t p = waitpid(-1, &st, rc != 0 && !killed ? WNOHANG : 0);
    if (p < 0) { if (errno == EINTR) continue; break; }
    if (p == 0) {  /* a rank failed, nobody else has exited yet */
      struct timespec now; clock_gettime(CLOCK_MONOTONIC, &now);
      double waited = (double)(now.tv_sec - t_fail.tv_sec) + 1e-9 * (double)(now.tv_nsec - t_fail.tv_nsec);
      if (waited > 0.25) {
        for (int r = 0; r < n; ++r) if (!done[r]) kill(pid[r], SIGTERM);
        killed = 1;
      } else {
        struct timespec nap = {0, 5 * 1000 * 1000}; nanosleep(&nap, NULL);
      }
      continue;
    }
    --left;
    for (int r = 0; r < n; ++r) if (pid[r] == p) done[r] = 1;
    int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st);
    if (code != 0 && rc == 0) { rc = code; clock_gettime(CLOCK_MONOTONIC, &t_fail); }
  }
  return rc;
}
