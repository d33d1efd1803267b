/* th_main.c -- `transit_b200 [-c file] [--key value ...]`: the reference's main()
 * (transit/src/transit.c:230-270): init, one spectrum, write it, clean up.
 *
 * TRANSIT_B200_GPUS=N (N = 2..8 GPUs of one box) shards the wavenumber range over N processes, one per GPU
 * (SURVEY 8e): this process forks N-1 more BEFORE any CUDA call, every rank reads the inputs, evaluates its own
 * band and joins the NCCL exchange inside libtransit_b200.so; all ranks end with the whole spectrum, rank 0
 * writes the files.  The output is byte-identical to a single-GPU run.                                     */
#include <stdio.h>
#include <stdlib.h>
#include <sys/wait.h>
#include <unistd.h>
#include "th_internal.h"

int main(int argc, char **argv) {
  const char *g = getenv("TRANSIT_B200_GPUS");
  int world = g ? atoi(g) : 1;
  if (world < 1) world = 1;
  if (world > 64) world = 64;
  pid_t kids[64];
  if (world > 1) {
    int rd[64];
    for (int r = 1; r < world; r++) {
      int fd[2];
      if (pipe(fd)) { perror("pipe"); return EXIT_FAILURE; }
      rd[r] = fd[0];
      th_idfd[r] = fd[1];
    }
    for (int r = 1; r < world; r++) {
      const pid_t p = fork();
      if (p < 0) { perror("fork"); return EXIT_FAILURE; }
      if (p == 0) {
        th_rank = r;
        th_idfd[r] = rd[r];
        if (!freopen("/dev/null", "w", stdout)) return EXIT_FAILURE;   /* rank 0 speaks for the run */
        break;
      }
      kids[r] = p;
    }
    th_world = world;
  }
  transit_init(argc, argv);
  double *out = (double *)malloc((size_t)(get_no_samples() > 0 ? get_no_samples() : 1) * sizeof(double));
  do_transit(out);
  if (th_rank == 0) th_write_spectrum();
  free_memory();
  free(out);
  if (th_rank == 0 && world > 1) {
    int bad = 0, st;
    for (int r = 1; r < world; r++)
      if (waitpid(kids[r], &st, 0) < 0 || !WIFEXITED(st) || WEXITSTATUS(st) != 0) bad = 1;
    if (bad) { fprintf(stderr, "a rank of the band-sharded run failed\n"); return EXIT_FAILURE; }
  }
  return EXIT_SUCCESS;
}
