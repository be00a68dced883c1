/*
 * zb200.h — C ABI of the B200-native Alard–Lupton image-subtraction path.
 *
 * This library replaces the `hotpants` executable that zuds fork/execs for every
 * image pair (zuds/subtraction.py:162, command line built by zuds/hotpants.py:77-93).
 * The reference crosses that boundary with FITS files and argv; here the same
 * inputs cross as plain pointers + a parameter struct.  No C++/torch types.
 *
 * Entry point            replaces (reference file:line)
 * ---------------------  ------------------------------------------------------
 * zb_params_default      hotpants/vargs.c:55-128      (defaults of every flag)
 * zb_params_check        hotpants/vargs.c:482-647     (FATAL ERROR sanity checks)
 * zb_sizes               hotpants/main.c:382-428,535-555 (derived sizes, region grid)
 * zb_subtract            hotpants/main.c:606-1867     (per-region loop: masks, stamp
 *                        search, fit, spatial_convolve, subtract, noise, output mask, QA)
 * zb_subtract_host       same, with host buffers (what subprocess.check_call +
 *                        FITS I/O deliver in the reference, zuds/subtraction.py:162)
 * zb_region_info etc.    per-region FITS keywords hotpants/main.c:1719-1819 and the
 *                        -hki kernel table main.c:573-579,1824-1835
 *
 * Conventions: images are row-major, x fastest, idx = x + nx*y (main.c:648-652);
 * float32 pixels; device pointers unless the name says host.  Return 0 on
 * success, negative zb_status on error (zb_last_error gives the text).  The
 * caller owns every buffer; a ctx owns only its private workspace and is used by
 * one host thread at a time.
 */
#ifndef ZB200_H
#define ZB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ZB_MAX_GAUSS   8
#define ZB_MAX_KSS     8     /* max -nss (sub-stamp centres kept per stamp) */

/* hotpants mask bits, hotpants/defaults.h:4-19 */
#define ZB_FLAG_BAD_PIXVAL   0x01
#define ZB_FLAG_SAT_PIXEL    0x02
#define ZB_FLAG_LOW_PIXEL    0x04
#define ZB_FLAG_ISNAN        0x08
#define ZB_FLAG_BAD_CONV     0x10
#define ZB_FLAG_INPUT_MASK   0x20
#define ZB_FLAG_OK_CONV      0x40
#define ZB_FLAG_INPUT_ISBAD  0x80
#define ZB_FLAG_T_BAD        0x100
#define ZB_FLAG_T_SKIP       0x200
#define ZB_FLAG_I_BAD        0x400
#define ZB_FLAG_I_SKIP       0x800
#define ZB_FLAG_OUTPUT_ISBAD 0x8000

typedef enum zb_status {
    ZB_OK              =  0,
    ZB_ERR_PARAM       = -1,   /* a vargs.c FATAL ERROR condition */
    ZB_ERR_UNSUPPORTED = -2,   /* valid hotpants option outside this build's scope */
    ZB_ERR_CUDA        = -3,
    ZB_ERR_NOMEM       = -4,
    ZB_ERR_INTERNAL    = -5
} zb_status;

/* mask element types accepted for -imi / -tmi */
#define ZB_MASK_U8   1
#define ZB_MASK_I16  2
#define ZB_MASK_I32  4

/* zb_region_info.flags */
#define ZB_REGION_SKIPPED   0x1   /* no usable stamps: region left untouched, main.c:1002-1009 */
#define ZB_REGION_SINGULAR  0x2   /* ludcmp saw an all-zero row, alard.c:1480-1493 (ignored, as in hotpants) */
#define ZB_REGION_RETRIED   0x4   /* stamp search redone with kerFitThresh*scaleFitThresh, main.c:948-995 */

/* Mirrors the hotpants globals that the path uses (hotpants/globals.h:27-60). */
typedef struct zb_params {
    int32_t nx, ny;                                   /* NAXIS1, NAXIS2 */
    int32_t hw_kernel, hw_ksstamp, n_ksstamps;        /* -r -rss -nss */
    int32_t ker_order, bg_order;                      /* -ko -bgo */
    int32_t nreg_x, nreg_y, nstamp_x, nstamp_y;       /* -nrx -nry -nsx -nsy */
    int32_t kc_step;                                  /* -kcs, 0 => fwKernel */
    int32_t ngauss;                                   /* -ng */
    int32_t deg[ZB_MAX_GAUSS];
    float   beta[ZB_MAX_GAUSS];                       /* 1/(2 sigma^2) as the float vargs.c:62-64 stores */
    float   t_uthresh, t_ukthresh, t_lthresh;         /* -tu -tuk -tl */
    float   i_uthresh, i_ukthresh, i_lthresh;         /* -iu -iuk -il */
    float   t_gain, t_rdnoise, i_gain, i_rdnoise;     /* -tg -tr -ig -ir */
    float   t_pedestal, i_pedestal;                   /* -tp -ip */
    float   ker_fit_thresh, scale_fit_thresh, min_frac_good_stamps;  /* -ft -sft -nft */
    float   stat_sig, ker_sig_reject, ker_frac_mask;  /* -ssig -ks -kfm */
    float   kf_spread_mask1, kf_spread_mask2;         /* -mins -mous */
    float   fill_val, fill_val_noise;                 /* -fi -fin */
    int32_t force_convolve;                           /* -c : 't' | 'i' | 'b' */
    int32_t phot_normalize;                           /* -n : 't' | 'i' | 'u' */
    int32_t fig_merit;                                /* -fom : 'v' | 's' | 'h' */
    int32_t convolve_variance;                        /* -convvar */
    int32_t rescale_ok;                               /* -okn */
    int32_t use_full_ss;                              /* (not a flag in 5.1.11; D_USEFULLSS) */
} zb_params;

/* Sizes derived exactly as hotpants/main.c:382-428 does. */
typedef struct zb_sizes {
    int32_t n_comp_ker, n_comp, n_bg_vectors, n_comp_total, mat_size;
    int32_t fw_kernel, fw_ksstamp, fw_stamp, kc_step;
    int32_t nstamp_x, nstamp_y, n_stamps;             /* per region (may be lowered, main.c:411-421) */
    int32_t n_regions, s_border;
    int32_t max_rpix_x, max_rpix_y;                   /* largest buffered region */
} zb_sizes;

/* Per-region results = the per-region FITS keywords of hotpants/main.c:1719-1819. */
typedef struct zb_region_info {
    int32_t rx_min, rx_max, ry_min, ry_max;           /* good-data bounds, 0-indexed inclusive (REGIONnn is these +1) */
    int32_t bx_min, bx_max, by_min, by_max;           /* buffered bounds */
    int32_t conv_template;                            /* CONVOLnn: 1 TEMPLATE, 0 IMAGE */
    int32_t flags;
    int32_t n_stamps;                                 /* stamps with >= 1 sub-stamp (compacted, main.c:896-904) */
    int32_t n_skipped;                                /* NskippedSubstamps of the last check_again */
    int32_t fit_iters;                                /* build/solve/check passes done */
    int32_t nx2norm;                                  /* NX2NRMnn */
    double  ksum;                                     /* KSUMnn */
    double  sssig, ssscat;                            /* SSSIGnn SSSCATnn */
    double  fsig, fscat;                              /* FSIGnn FSCATnn */
    double  x2norm;                                   /* X2NRMnn */
    double  dmean, dsige, dsig;                       /* DMEANnn DSIGEnn DSIGnn */
    double  dmeano, dsigeo, dsigo;                    /* DMEANOnn DSIGEOnn DSIGOnn */
    double  nscalo;                                   /* NSCALOnn (sd/nmean) */
    double  t_merit, i_merit;                         /* -c b only: check_stamps' merit of each direction (alard.c:483-724;
                                                         the "Result : merit" lines of main.c:1037,1053), else 0 */
} zb_region_info;

/* Per-stamp bookkeeping = hotpants/globals.h:1-24 minus the big arrays (what -savexy dumps). */
typedef struct zb_stamp_info {
    int32_t x0, y0;                                   /* stamp origin, region coords */
    int32_t nss, sscnt;                               /* sub-stamps found / index in use (== nss: exhausted) */
    int32_t xss[ZB_MAX_KSS], yss[ZB_MAX_KSS];         /* sub-stamp centres, region coords */
    double  mode, fwhm;                               /* sky and sky width from getStampStats3 */
    double  chi2;                                     /* last figure of merit */
} zb_stamp_info;

typedef struct zb_ctx zb_ctx;

const char *zb_version(void);
void  zb_params_default(zb_params *p);
int   zb_params_check(const zb_params *p, char *msg, int msglen);
int   zb_sizes_of(const zb_params *p, zb_sizes *s);

int   zb_create(int device, zb_ctx **out);
void  zb_destroy(zb_ctx *ctx);
const char *zb_last_error(const zb_ctx *ctx);

/*
 * One image pair, everything resident on the ctx's device.
 *   sci, ref            float32 [ny*nx]                      (-inim, -tmplim)
 *   sci_rms, ref_rms    float32 [ny*nx] or NULL              (-ini, -tni; NULL => gain/rdnoise model, functions.c:1172)
 *   sci_mask, ref_mask  [ny*nx] of *_dtype or NULL           (-imi, -tmi; >0 means masked, main.c:778-803)
 *   diff, noise         float32 [ny*nx] outputs              (-outim, -oni)
 *   out_mask            int32 [ny*nx] output or NULL         (-omi)
 *   regions             host, [n_regions]                    out
 *   kernel_sol          host, [n_regions*(n_comp_total+1)] or NULL   (-hki table column per region)
 *   stamps              host, [n_regions*n_stamps] or NULL
 *   stream              cudaStream_t (as void*).  NULL has CUDA's own meaning, the legacy default stream, so
 *                       work a caller queued on its default stream (torch's, say) is ordered before this call;
 *                       ZB_STREAM_OWN selects the context's private non-blocking stream (what batch drivers
 *                       that run many contexts side by side want).
 * Returns when the results are complete (host outputs are filled, device outputs
 * are ordered on `stream`).
 */
#define ZB_STREAM_OWN ((void *)(intptr_t)-1)
int zb_subtract(zb_ctx *ctx, const zb_params *p,
                const float *sci, const float *ref,
                const float *sci_rms, const float *ref_rms,
                const void *sci_mask, int sci_mask_dtype,
                const void *ref_mask, int ref_mask_dtype,
                float *diff, float *noise, int32_t *out_mask,
                zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps,
                void *stream);

/* Same call with HOST pointers for every array: copies in, runs, copies out. */
int zb_subtract_host(zb_ctx *ctx, const zb_params *p,
                     const float *sci, const float *ref,
                     const float *sci_rms, const float *ref_rms,
                     const void *sci_mask, int sci_mask_dtype,
                     const void *ref_mask, int ref_mask_dtype,
                     float *diff, float *noise, int32_t *out_mask,
                     zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps);

/* npairs subtractions from HOST buffers, software-pipelined inside the context: the inputs of pair
 * n+1 are copied to the device and the products of pair n-1 back to the host while pair n is computed; one
 * cudaMemcpyAsync per array, pinned caller buffers make them asynchronous.  What a batch driver in any language calls
 * instead of running many contexts side by side for the overlap (zuds: one hotpants process per pair, srun -n 64).
 *   params   [npairs]                         one parameter set per pair (same frame size; thresholds usually differ)
 *   pairs    [npairs]                         host pointers of each pair (out_mask NULL: the int mask stays on the device,
 *                                             as with a zuds command line, which has no -omi)
 *   regions  [npairs * n_regions], kernel_sol [npairs * n_regions * (n_comp_total+1)] or NULL, stamps [...] or NULL
 *   rcs      [npairs] per-pair status, or NULL.  Returns the first non-OK status after the whole batch has run. */
typedef struct zb_host_pair {
    const float *sci, *ref, *sci_rms, *ref_rms;
    const void  *sci_mask, *ref_mask;
    int32_t      sci_mask_dtype, ref_mask_dtype;
    float       *diff, *noise;
    int32_t     *out_mask;
} zb_host_pair;
int zb_subtract_batch(zb_ctx *ctx, int npairs, const zb_params *params, const zb_host_pair *pairs,
                      zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps, int *rcs);

/* Timing/diagnostics of the last zb_subtract on this ctx (ms per stage, CUDA events). */
#define ZB_NSTAGES 12
int zb_last_timings(const zb_ctx *ctx, float *ms, int n);
const char *zb_stage_name(int i);
/* number of kernel launches issued by this ctx since creation */
long long zb_launch_count(const zb_ctx *ctx);
/* sizeof() of the public structs as built (0 zb_params, 1 zb_sizes, 2 zb_region_info, 3 zb_stamp_info, 4 zb_tile_info, 5 zb_host_pair):
   lets a binding verify its own struct mirror before the first call */
int zb_abi_sizeof(int which);

/* ------------------------------------------------------------------------------------------
 * Wrapper pre-step on the GPU: zuds.utils.quick_background_estimate, zuds/utils.py:32-53.
 *   bkg = median(data[mask == 0]);  bkgstd = 1.4826 * median(|data[mask == 0] - bkg|)
 * evaluated as numpy does on float32 pixels: float32 medians (the two central values of an even count are added
 * and halved in float32), float32 deviations, the factor 1.4826 in double; NaN when nothing is unmasked or an
 * unmasked pixel is NaN.  Exact order statistics (radix selection),
 * so the -il/-tl thresholds zuds derives from them are identical to the host computation.
 * mask may be NULL (every pixel used); n = number of pixels. */
int zb_background_estimate(zb_ctx *ctx, const float *data /* device */, const void *mask /* device */, int mask_dtype,
                           long long n, double *bkg, double *bkgstd, void *stream);
int zb_background_estimate_host(zb_ctx *ctx, const float *data /* host */, const void *mask /* host */, int mask_dtype,
                                long long n, double *bkg, double *bkgstd);

/* The same estimate for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * It takes no context and calls no cudaMalloc / cudaFree / cudaMallocHost or any allocating library routine: the
 * selection state (histograms, prefixes, counts) lives in a static device array of ZB_BKG_STATIC_SCRATCH_BYTES that
 * every GPU loading the library holds until exit.  Runs on the calling thread's CURRENT device; does not synchronise.
 *   data, mask   device pointers as for zb_background_estimate (mask may be NULL)
 *   n            number of pixels, at most ZB_BKG_STATIC_MAX_PIXELS (the scratch's histogram counters are 32-bit);
 *                a larger n is refused with ZB_ERR_PARAM and a message, before anything is launched
 *   out          DEVICE buffer of two doubles: out[0] = bkg, out[1] = bkgstd, complete when `stream` reaches this point
 *   stream       a cudaStream_t, or NULL for the legacy default stream (ZB_STREAM_OWN is refused: there is no context)
 *   msg, msglen  optional buffer for the error text (what zb_last_error is for the context calls)
 * All calls on a device share the one array, and a call finds the previous call's contents in it: every call clears
 * the state on `stream` first, and calls that OVERLAP on the GPU (streams not ordered with each other) race — order
 * them.  The first call on a device resolves the array's device address and loads the kernels; make it outside
 * graph capture. */
#define ZB_BKG_STATIC_SCRATCH_BYTES 17408
#define ZB_BKG_STATIC_MAX_PIXELS    (1LL << 31)
int zb_background_estimate_static(const float *data /* device */, const void *mask /* device */, int mask_dtype, long long n,
                                  double *out /* device [2] */, void *stream, char *msg, int msglen);

/* ------------------------------------------------------------------------------------------
 * The other elementwise steps zuds does around the hotpants call, on DEVICE pointers so that a
 * caller holding its images in HBM never returns to the host between them, zb_background_estimate and zb_subtract.
 * All bit-identical to the numpy statements they replace (integer work / correctly rounded float32).
 *   zb_mask_combine        badpix = refmask | scimask (zuds/subtraction.py:135) into out_or (int32, may be NULL) and
 *                          its boolean (badpix & bad_sum) > 0 (zuds/mask.py:42-72; bad_sum = 202749 for ZTF) into
 *                          out_bool (uint8 0/1, what -imi reads; may be NULL).  Either input mask may be NULL (= 0).
 *   zb_mask_flag_sentinel  mask |= bit_value where diff == sentinel, exact float equality
 *                          (zuds/subtraction.py:167-177: sentinel 1e-30f, bit_value 2**17)
 *   zb_rms_from_weight     rms = 1/sqrt(weight); big_rms where bad[i] != 0 or data[i] >= sat_level (zuds/image.py:179-190;
 *                          sat_level = (float)(0.9 * SATURATE); bad and data may be NULL)
 *   zb_weight_from_rms     weight = 1/rms^2; 0 where bad[i] != 0 or data[i] >= sat_level (zuds/image.py:146-149)
 * n = number of pixels; `stream` as in zb_subtract; the calls do not synchronise. */
int zb_mask_combine(zb_ctx *ctx, const void *sci_mask, int sci_dtype, const void *ref_mask, int ref_dtype, long long n,
                    int bad_sum, int32_t *out_or, uint8_t *out_bool, void *stream);
int zb_mask_flag_sentinel(zb_ctx *ctx, int32_t *mask, const float *diff, long long n, float sentinel, int bit_value, void *stream);
int zb_rms_from_weight(zb_ctx *ctx, const float *weight, const uint8_t *bad, const float *data, float sat_level, float big_rms,
                       long long n, float *rms, void *stream);
int zb_weight_from_rms(zb_ctx *ctx, const float *rms, const uint8_t *bad, const float *data, float sat_level,
                       long long n, float *weight, void *stream);

/* ------------------------------------------------------------------------------------------
 * Alignment of the reference onto the science grid, the step before the subtraction: what zuds runs
 * SWarp for, once per image and once per mask (zuds/swarp.py:107-204, zuds/fitsfile.py:290-314, configuration
 * astromatic/makecoadd/default.swarp: LANCZOS3, no oversampling, fixed flux scale, one input image).
 *   zb_align_lanczos3  out[j*dnx+i] = 6x6 Lanczos-3 interpolation of `src` at the position of target pixel (i, j),
 *                      weights of each axis normalised to unit sum, times |det CD_dst| / |det CD_src|;
 *                      value 0 and weight 0 where the 6x6 footprint leaves the source frame, weight 1 elsewhere
 *                      (out_weight may be NULL).  Device pointers.
 *   zb_align_mask      nearest-neighbour transport of an integer mask (flag bits cannot be interpolated), with
 *                      `nodata_bit` OR-ed in where the image's weight is 0 (zuds/mask.py:26-33: 2**16).
 * World coordinates: gnomonic (TAN) projection, FITS conventions (CRPIX 1-based, CRVAL / CD in degrees).  Distortion
 * polynomials (TPV PVi_j) are not built. */
typedef struct zb_wcs {
    double crpix1, crpix2, crval1, crval2;
    double cd11, cd12, cd21, cd22;
} zb_wcs;
int zb_align_lanczos3(zb_ctx *ctx, const float *src, int snx, int sny, const zb_wcs *src_wcs,
                      int dnx, int dny, const zb_wcs *dst_wcs, float *out, float *out_weight, void *stream);
int zb_align_mask(zb_ctx *ctx, const void *src_mask, int dtype, int snx, int sny, const zb_wcs *src_wcs,
                  int dnx, int dny, const zb_wcs *dst_wcs, int32_t *out, int nodata_bit, void *stream);

/* ------------------------------------------------------------------------------------------
 * Row-tile sessions: ONE oversized frame fitted by a single region (nreg_x == nreg_y == 1,
 * BASELINE.json configs[4]) split over `world` GPUs by image rows, one process (rank) per GPU.
 * The reference has no counterpart (one single-threaded process per frame, zuds/mpi.py:36-64);
 * the arithmetic per pixel / per stamp is that of zb_subtract and the results are bit-identical
 * to zb_subtract on the whole frame, whatever `world` is.
 *
 * Every rank keeps full-size [ny*nx] device arrays but only needs rows [in_row0, in_row1) of the
 * inputs: it loads its own rows [row0, row1) and receives the halo rows from its neighbours
 * (P2P exchange, done by the caller).  Stamp-grid rows [cell_row0, cell_row1) and convolution
 * block rows [blk_row0, blk_row1) are the rank's; it writes output rows [out_row0, out_row1).
 * Data crossing ranks, all summed with zeros elsewhere so that one all-reduce(SUM) merges them bit
 * for bit: the stamp-search records, the per-stamp normal equations and the per-stamp figures of
 * merit.  The global normal matrix is then built and solved on every rank from the same per-stamp
 * terms in the reference's stamp order (alard.c:843-1025), so no rank-dependent summation order
 * enters the solution.  Mask rows changed by the stamp search travel with a second halo exchange.
 *
 *   zb_tile_plan          pure host arithmetic: who owns what (no GPU needed)
 *   zb_tile_begin         configure, input masks of the needed rows, stamp search of the own grid cells
 *   zb_tile_cells_export/import   search records as int32 words, zero outside the rank's cells
 *   zb_tile_resolve       replicated on every rank: stale-slot fix-ups, threshold retry, stamp compaction.
 *                         *action: 0 done | 1 cell *cell must be searched again: every rank calls
 *                         zb_tile_research(cell) (a no-op except on its owner), then the records are
 *                         merged again | 2 every rank calls zb_tile_begin_retry, then merges again
 *   zb_tile_rows          copy rows [r0, r1) of a per-region work array out of / into the context:
 *                         the input-stage mask (after resolve, for the mask halo exchange) and the
 *                         pre-output-masking difference / noise / mask the QA statistics read
 *   zb_tile_fill          fill the rank's own stamps (vectors, per-stamp normal equations)
 *   zb_tile_normal_export/import  per-stamp Gram matrices + right-hand sides (device doubles)
 *   zb_tile_solve         replicated global fit; figure of merit of the rank's own stamps
 *   zb_tile_sig_export/import     per-stamp figure of merit (device doubles)
 *   zb_tile_clip          replicated check_again; refills own replaced stamps; *active: another pass
 *   zb_tile_convolve      the rank's block rows of the fused convolution/output pass
 *   zb_tile_finish        rank 0, after the rows were gathered: FSIG, region statistics, records
 * Every call runs on the stream given to zb_tile_begin (pass the stream the collectives use).
 */
typedef struct zb_tile_info {
    int32_t rank, world;
    int32_t cell_row0, cell_row1;     /* stamp-grid rows */
    int32_t row0, row1;               /* image rows this rank loads / is authoritative for */
    int32_t in_row0, in_row1;         /* image rows it needs (own + halo) */
    int32_t blk_row0, blk_row1;       /* convolution block rows */
    int32_t out_row0, out_row1;       /* output rows it writes */
    int32_t halo;                     /* max(row0 - in_row0, in_row1 - row1) */
    int32_t n_cells, cell_words;      /* grid cells of the frame, int32 words per search record */
    int32_t n_slots, gram_len, rhs_len;  /* stamp slots; doubles per slot in the exported normal equations */
} zb_tile_info;

enum { ZB_TILE_MASK_IN = 0, ZB_TILE_RDIFF = 1, ZB_TILE_RNOISE = 2, ZB_TILE_OMASK = 3 };

int zb_tile_plan(const zb_params *p, int world, int rank, zb_tile_info *out);
int zb_tile_begin(zb_ctx *ctx, const zb_params *p, int world, int rank,
                  const float *sci, const float *ref, const float *sci_rms, const float *ref_rms,
                  const void *sci_mask, int sci_mask_dtype, const void *ref_mask, int ref_mask_dtype,
                  float *diff, float *noise, int32_t *out_mask, void *stream);
int zb_tile_begin_retry(zb_ctx *ctx);
int zb_tile_cells_export(zb_ctx *ctx, int32_t *words /* host [n_cells*cell_words] */);
int zb_tile_cells_import(zb_ctx *ctx, const int32_t *words);
int zb_tile_resolve(zb_ctx *ctx, int *action, int *cell);
int zb_tile_research(zb_ctx *ctx, int cell);
int zb_tile_rows(zb_ctx *ctx, int which, int import, int r0, int r1, void *buf /* device [(r1-r0)*nx] 4-byte elements */);
int zb_tile_fill(zb_ctx *ctx);
int zb_tile_normal_export(zb_ctx *ctx, double *gram /* device [n_slots*gram_len] */, double *rhs /* device [n_slots*rhs_len] */);
int zb_tile_normal_import(zb_ctx *ctx, const double *gram, const double *rhs);
int zb_tile_solve(zb_ctx *ctx);
int zb_tile_sig_export(zb_ctx *ctx, double *sig /* device [n_slots] */);
int zb_tile_sig_import(zb_ctx *ctx, const double *sig);
int zb_tile_clip(zb_ctx *ctx, int *active);
int zb_tile_convolve(zb_ctx *ctx);
int zb_tile_finish(zb_ctx *ctx, zb_region_info *regions, double *kernel_sol, zb_stamp_info *stamps);

#ifdef __cplusplus
}
#endif
#endif /* ZB200_H */
