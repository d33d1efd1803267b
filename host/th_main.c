/* th_main.c -- `transit_b200 [-c file] [--key value ...]`: the reference's main()
 * (transit/src/transit.c:230-270): init, one spectrum, write it, clean up. */
#include <stdlib.h>
#include "th_internal.h"

int main(int argc, char **argv) {
  transit_init(argc, argv);
  double *out = (double *)malloc((size_t)(get_no_samples() > 0 ? get_no_samples() : 1) * sizeof(double));
  do_transit(out);
  th_write_spectrum();
  free_memory();
  free(out);
  return EXIT_SUCCESS;
}
