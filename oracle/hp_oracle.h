/*
 * hp_oracle.h — CPU restatement of the reference hotpants path (TEST INFRASTRUCTURE).
 *
 * This is the checker the CUDA path is compared against; it is never the thing
 * shipped or measured.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg may load it.  Parity of this restatement is PINNED against
 * the reference itself: oracle/_ref/hotpants (the unmodified reference sources,
 * built by oracle/Makefile) run on the same synthetic inputs — see
 * tests/test_oracle_vs_ref.py and tests/golden/.
 *
 * Structs are shared with the product's public header so both sides can be
 * driven by the same Python ctypes definitions.
 */
#ifndef HP_ORACLE_H
#define HP_ORACLE_H

#include "../include/zb200.h"

#ifdef __cplusplus
extern "C" {
#endif

void hp_oracle_params_default(zb_params *p);
int  hp_oracle_sizes(const zb_params *p, zb_sizes *s);

/* host pointers; masks are int32 (>0 = masked) or NULL; outputs caller-allocated */
int hp_oracle_subtract(const zb_params *p,
                       const float *sci, const float *ref,
                       const float *sci_rms, const float *ref_rms,
                       const int32_t *sci_mask, const int32_t *ref_mask,
                       float *diff, float *noise, int32_t *out_mask,
                       zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps);

/* building blocks exported for unit tests */
double hp_oracle_ran1(int *idum);
int    hp_oracle_sigma_clip(const float *data, int count, double *mean, double *stdev, int maxiter, float stat_sig);
void   hp_oracle_basis(const zb_params *p, double *kernel_vec /* [nCompKer][fw*fw] */,
                       double *filter_x, double *filter_y /* [nCompKer][fw] */);

#ifdef __cplusplus
}
#endif
#endif
