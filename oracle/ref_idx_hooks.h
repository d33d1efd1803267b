/* oracle/ref_idx_hooks.h -- TEST INFRASTRUCTURE ONLY.  Force-included (gcc -include) into the patched
 * copy of the reference's extinction.c that oracle/Makefile generates under oracle/_ref/obj: the three
 * calls sed adds there (extinction.c:449,467-470,496) land in these hooks, defined in ref_dump.c. */
#ifndef REF_IDX_HOOKS_H
#define REF_IDX_HOOKS_H
extern long ref_idx_lead;
void ref_idx_hook(long lead, long last, long iown, long idwn, int idop, int ilor, long minj, long maxj);
void ref_idx_skip(long lead, long last, long iown);
#endif
