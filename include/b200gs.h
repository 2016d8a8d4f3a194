/*
 * b200gs.h -- C ABI of libb200gs.so, the sm_100a implementation of the
 * differentiable rendering hot path of gssales/gaussian-splatting-reflection.
 *
 * Every entry point takes plain device pointers, sizes and a CUDA stream
 * (passed as void*; NULL = legacy default stream, which is what the reference
 * launches on).  No torch types cross this boundary.  All functions return 0 on
 * success and a negative B200GS_E* code on failure; b200gs_last_error() gives
 * the message of the calling thread's last failure.
 *
 * Reference interfaces replaced (paths relative to the reference's
 * submodules/):
 *   surfel:  diff-surfel-rasterization/cuda_rasterizer/rasterizer.h:20-94
 *            (CudaRasterizer::Rasterizer::{markVisible,forward,backward}) as
 *            called from rasterize_points.cu:39-288
 *   3DGS:    diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:20-102
 *   cubemap: cubemapencoder/src/cubemapencoder.h:6-17
 *
 * Scratch ("geom", "img", "binning") is caller-allocated device memory whose
 * layout is private to this library (it is NOT the reference's fromChunk
 * layout); the *_bytes functions give the sizes and the *_inspect_* functions
 * unpack it into the reference's SoA arrays for parity checks.
 */
#ifndef B200GS_H_INCLUDED
#define B200GS_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200GS_OK 0
#define B200GS_EINVAL (-1)   /* bad argument (message says which)            */
#define B200GS_ECUDA (-2)    /* CUDA runtime error (message = cudaGetErrorString) */
#define B200GS_ESIZE (-3)    /* scratch buffer too small                      */

const char* b200gs_last_error(void);
/* Library/ABI version and the SM architecture the kernels were built for. */
int b200gs_version(void);
int b200gs_sm_arch(void);
/* Number of kernel launches this library has issued since load (all threads);
 * bench.py reads it around the timed region for its "gpu_launches" claim.    */
unsigned long long b200gs_launch_count(void);
/* sizeof() of the argument structs of this header as the library was compiled, for bindings to check their own
 * layout against: 0 b200gs_surfel_fwd, 1 b200gs_surfel_bwd, 2 b200gs_gauss_fwd, 3 b200gs_gauss_bwd, 4 b200gs_adam,
 * 5 b200gs_ply_layout; 0 for any other index.                                                                  */
size_t b200gs_struct_bytes(int which);

/* Per-stage device timing for bench.py: when enabled, every stage (kernel) is
 * bracketed by a CUDA event pair on the stream it is launched on.
 * b200gs_timing_read waits for the recorded events, adds the elapsed ms and the
 * launch counts per stage into ms[0..n) / counts[0..n) and clears the record.   */
void b200gs_timing_enable(int on);
/* Events around ONE stage only (index as in b200gs_timing_stage_name): an event pair between every two
 * kernels keeps the next launch from overlapping the previous kernel's tail, which costs ~6 % of a step
 * when every stage is instrumented; one pair does not.  stage < 0: off. */
void b200gs_timing_only(int stage);
int b200gs_timing_stages(void);
const char* b200gs_timing_stage_name(int stage);
int b200gs_timing_read(float* ms, unsigned long long* counts, int n);

/* Instance lists (field `exact_lists` of b200gs_surfel_fwd / b200gs_gauss_fwd; a per-call argument, the
 * library keeps no mode of its own).  0: culled emission -- a (Gaussian, tile) instance is
 * only emitted if the Gaussian can reach alpha >= 1/255 somewhere in the tile, so num_rendered and
 * the sorted lists are a subsequence of the reference's (typically a few per cent of it when splats
 * cross the camera plane); colour/aux/reflection images, radii, weights and gradients are
 * unchanged.  1: exact lists -- num_rendered, the sorted (key, value) lists, the tile ranges and
 * n_contrib are the reference's, bit for bit (S/cuda_rasterizer/rasterizer_impl.cu:282-321); the
 * compositing kernels then skip the unreachable instances through the compacted list instead.
 * Stage 1 and stage 2 of one forward must be given the same struct.                          */

/* Cull audit (test infrastructure for the sub-tile culling of the surfel compositing; field `cull_audit`
 * of b200gs_surfel_fwd, to be combined with exact_lists = 1): the forward compositing ignores the
 * per-instance hit masks, evaluates every
 * (warp block, Gaussian) pair the way the reference does and counts the pairs that pass the
 * reference's alpha >= 1/255 cut although their mask bit is clear -- out2[0] with the forward's
 * `unstable` threshold (S/cuda_rasterizer/forward.cu:373), out2[1] with the backward's
 * (backward.cu:302).  Both must read 0: the masks may only drop pairs the reference drops.
 * The counters live on the device the call ran on.                                          */
int b200gs_debug_cull_violations(unsigned long long* out2, int reset);
/* Workload counters of the audit pass: out8[0] (warp block, entry) pairs visited with the hit
 * bit set, [1] those with >= 1 contributing pixel, [2] contributing (pixel, entry) pairs,
 * [3] sum over (warp, batch) of the busiest pixel's contributions, [4] (warp, batch) count.   */
int b200gs_debug_cull_stats(unsigned long long* out8, int reset);

/* ------------------------------------------------------------------------ */
/* markVisible  (S/rasterize_points.cu:269-288, rasterizer_impl.cu:54-66)     */
int b200gs_mark_visible(int P, const float* means3D, const float* viewmatrix,
                        const float* projmatrix, uint8_t* present, void* stream);

/* ------------------------------------------------------------------------ */
/* Surfel (2DGS) rasterizer                                                   */

/* Inputs of one forward view; mirrors the argument list of
 * CudaRasterizer::Rasterizer::forward (S/cuda_rasterizer/rasterizer_impl.cu:198-225).
 * Optional pointers are NULL when "not provided" (the reference tests
 * data_ptr()==nullptr of an empty tensor).                                   */
typedef struct b200gs_surfel_fwd {
  int P;                 /* #Gaussians                                        */
  int sh_degree;         /* D: active SH degree 0..3                          */
  int sh_coeffs;         /* M: coefficients per channel stored in shs (<=16)  */
  int width, height;
  float tan_fovx, tan_fovy;
  float scale_modifier;
  int prefiltered;       /* !=0: a culled point is an error (B200GS_EINVAL after the stage) */
  int debug;             /* !=0: synchronise + check after every stage        */
  const float* background;        /* [3]                                      */
  const float* means3D;           /* [P,3]                                    */
  const uint8_t* env_scope_mask;  /* bool, indexed by Gaussian id (>=P entries) or NULL */
  const float* shs;               /* [P,M,3] or NULL                          */
  const float* colors_precomp;    /* [P,3]  or NULL                           */
  const float* refl_strengths;    /* [P]    or NULL (=0)                      */
  const float* opacities;         /* [P]                                      */
  const float* scales;            /* [P,2]  or NULL                           */
  const float* rotations;         /* [P,4] (w,x,y,z) or NULL                  */
  const float* transMat_precomp;  /* [P,9]  or NULL                           */
  const float* viewmatrix;        /* [16]                                     */
  const float* projmatrix;        /* [16]                                     */
  const float* campos;            /* [3]                                      */
  void* geom;   size_t geom_bytes;     /* >= b200gs_surfel_geom_bytes(P)      */
  void* img;    size_t img_bytes;      /* >= b200gs_surfel_image_bytes(W,H)   */
  void* binning; size_t binning_bytes; /* >= b200gs_binning_bytes(R); stage 2 */
  float* out_color;        /* [3,H,W]                                         */
  float* out_others;       /* [8,H,W] depth,alpha,normal xyz,median,distortion,mask */
  float* out_refl;         /* [1,H,W]                                         */
  int* radii;              /* [P]                                             */
  float* gaussian_weights; /* [P], zeroed by stage 1, maximised by stage 2    */
  int exact_lists;         /* instance-list mode of this call, see above      */
  int cull_audit;          /* != 0: audit pass (test infrastructure), see above */
  int depth_slices;        /* culled emission only.  0: automatic; 1: one slice (no occlusion-aware binning);
                            * n <= B200GS_MAX_SLICES: at most n depth slices -- see b200gs_surfel_forward_stage1 */
  float slice_frac[7];     /* with depth_slices = n > 1 and slice_frac[0] > 0: slice k ends where the cumulated instance
                            * count reaches slice_frac[k] * num_rendered (ascending, < 1; n - 1 values); otherwise the
                            * library's schedule (0.05, 0.25)                                                        */
} b200gs_surfel_fwd;

size_t b200gs_surfel_geom_bytes(int P);
size_t b200gs_surfel_image_bytes(int width, int height);
size_t b200gs_binning_bytes(int num_rendered);

/* Stage 1: preprocess, reachable-tile counts, depth order, prefix sum, slice plan.  Writes the plan of the forward
 * to plan[0 .. B200GS_PLAN_INTS) (host memory) and returns after the stream has drained -- the same single host
 * sync the reference has at rasterizer_impl.cu:285-286.  plan[0] = num_rendered (the total number of
 * (Gaussian, tile) instances: size `binning` for it), plan[1] = S, the number of depth slices, then the first
 * Gaussian (in depth order) and the first instance of every slice.  Needs geom + img only.
 * Depth slices (culled emission): the instance list is built and composited front to back in S pieces; tiles whose
 * pixels have all reached the reference's T < 1e-4 stop (forward.cu:430-434) receive no instances from later pieces,
 * which are then not emitted, sorted or tested at all.  The blending order within a tile is unchanged, so the
 * images, radii, weights and gradients are the same bits as with one slice.                                       */
#define B200GS_MAX_SLICES 8
#define B200GS_PLAN_INTS 24
int b200gs_surfel_forward_stage1(const b200gs_surfel_fwd* a, int* plan, void* stream);
/* Stage 2: per slice duplicate-with-keys, tile sort, tile ranges, block masks, per-tile compositing.  `plan` as
 * written by stage 1; needs binning sized for plan[0].  Asynchronous.         */
int b200gs_surfel_forward_stage2(const b200gs_surfel_fwd* a, const int* plan, void* stream);

/* Backward; mirrors CudaRasterizer::Rasterizer::backward
 * (S/cuda_rasterizer/rasterizer_impl.cu:358-466).  All dL_d* outputs must be
 * zero-filled by the caller (the reference shim does torch::zeros,
 * S/rasterize_points.cu:207-226).  `accum` is library scratch.               */
typedef struct b200gs_surfel_bwd {
  int P, sh_degree, sh_coeffs, num_rendered;
  int width, height;
  float tan_fovx, tan_fovy, scale_modifier;
  int debug;
  const float* background;
  const float* means3D;
  const float* shs;               /* or NULL */
  const float* colors_precomp;    /* or NULL */
  const float* scales;            /* or NULL */
  const float* rotations;         /* or NULL */
  const float* transMat_precomp;  /* or NULL */
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  const int* radii;
  const void* geom; const void* binning; const void* img;  /* from the forward */
  const float* dL_dout_color;   /* [3,H,W] */
  const float* dL_dout_others;  /* [8,H,W] */
  const float* dL_dout_refl;    /* [1,H,W] or NULL (=0; the reference dereferences it unconditionally) */
  void* accum; size_t accum_bytes;  /* >= b200gs_surfel_accum_bytes(P); need not be zeroed */
  float* dL_dmeans2D;    /* [P,3] */
  float* dL_dnormal;     /* [P,3] */
  float* dL_dopacity;    /* [P]   */
  float* dL_dcolors;     /* [P,3] */
  float* dL_drefl;       /* [P] or NULL */
  float* dL_dmeans3D;    /* [P,3] */
  float* dL_dtransMat;   /* [P,9] */
  float* dL_dsh;         /* [P,M,3] or NULL */
  float* dL_dscales;     /* [P,2] */
  float* dL_drotations;  /* [P,4] */
  /* != 0: the parameter gradients (dL_dopacity, dL_drefl, dL_dmeans3D, dL_dsh, dL_dscales,
   * dL_drotations) are ADDED to the buffers' contents for the Gaussians this view reached and left
   * alone for the others, so the gradients of a batch of views accumulate in place (no reference
   * counterpart: autograd does it there with one add per tensor per view).  0: overwritten; rows
   * of culled Gaussians are not written in either mode (callers zero-fill, S/rasterize_points.cu:186-197). */
  int accumulate;
  /* sizes of the forward's scratch as handed back (checked against what P, width x height and num_rendered need;
   * the reference re-derives pointers from the raw bytes unchecked, rasterizer_impl.cu:393-395) */
  size_t geom_bytes, binning_bytes, img_bytes;
} b200gs_surfel_bwd;

size_t b200gs_surfel_accum_bytes(int P);
int b200gs_surfel_backward(const b200gs_surfel_bwd* a, void* stream);

/* The same backward for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.
 * `a->accum` / `a->accum_bytes` are IGNORED; the accumulator lives in a file-scope __device__ array inside the
 * library (csrc/static_scratch.cu), sized for at most B200GS_STATIC_ACCUM_MAX_POINTS Gaussians (10 MiB, held by
 * every GPU that loads the library until the process exits).  P beyond that maximum -> B200GS_ESIZE with a message.
 * Calls no cudaMalloc / cudaFree / allocating routine.  All calls on a device share the one array (the rows a call
 * uses are cleared on its stream first), so calls that overlap on the GPU race: serialise them.  Inputs, outputs
 * and everything else are as for b200gs_surfel_backward.                                                        */
#define B200GS_STATIC_ACCUM_MAX_POINTS 131072
int b200gs_surfel_backward_static(const b200gs_surfel_bwd* a, void* stream);

/* Parity-check helpers: unpack the private scratch layouts into the SoA arrays
 * the reference keeps in GeometryState / BinningState / ImageState
 * (S/cuda_rasterizer/rasterizer_impl.h:33-69).  Any output may be NULL.      */
int b200gs_surfel_inspect_geom(const void* geom, int P,
                               float* depths, float* means2D /*[P,2]*/, float* transMat /*[P,9]*/,
                               float* normal_opacity /*[P,4]*/, float* rgb /*[P,3]*/,
                               uint32_t* tiles_touched, uint32_t* point_offsets,
                               uint8_t* clamped /*[P,3]*/, void* stream);
/* The library sorts 32-bit (tile, id) pairs emitted in depth order; the reference's 64-bit keys
 * (tile << 32 | depth bits) are rebuilt here from `depths` (float[P] as returned by
 * *_inspect_geom; needed only when point_list_keys != NULL).                              */
int b200gs_inspect_binning(const void* binning, int num_rendered, const float* depths,
                           uint64_t* point_list_keys, uint32_t* point_list, void* stream);
int b200gs_surfel_inspect_image(const void* img, int width, int height,
                                float* accum_alpha /*[3,N]*/, uint32_t* n_contrib /*[2,N]*/,
                                uint32_t* ranges /*[T,2]*/, void* stream);

/* Depth-slice bookkeeping of the last forward that used this img scratch (meaningful when it had more than one
 * slice): tile_done [tiles] = 1 where every pixel of the tile saturated before the last slice; live [1] = tiles
 * that did not.  Either may be NULL.                                                                          */
int b200gs_surfel_inspect_tiles(const void* img, int width, int height, uint8_t* tile_done, uint32_t* live, void* stream);

/* ------------------------------------------------------------------------ */
/* 3DGS (EWA) rasterizer with normal / reflection-strength / inverse-depth channels.
 * Mirrors CudaRasterizer::Rasterizer::{forward,backward} of the reference's
 * diff-gaussian-rasterization (G/cuda_rasterizer/rasterizer_impl.cu:198-356, :358-472) as
 * called from G/rasterize_points.cu:39-286.  Same two-stage forward as the surfel path.     */
typedef struct b200gs_gauss_fwd {
  int P, sh_degree, sh_coeffs, width, height;
  float tan_fovx, tan_fovy, scale_modifier;
  int prefiltered, antialiasing, debug;
  const float* background;        /* [3]                                      */
  const float* means3D;           /* [P,3]                                    */
  const float* shs;               /* [P,M,3] or NULL                          */
  const float* colors_precomp;    /* [P,3]  or NULL                           */
  const float* normals;           /* [P,3]  per-Gaussian normals, composited as given */
  const float* refl_strengths;    /* [P]                                      */
  const float* opacities;         /* [P]                                      */
  const float* scales;            /* [P,3]  or NULL                           */
  const float* rotations;         /* [P,4] (w,x,y,z), NOT normalised by the kernels (G/forward.cu:123) */
  const float* cov3D_precomp;     /* [P,6]  or NULL                           */
  const float* viewmatrix;        /* [16] */
  const float* projmatrix;        /* [16] */
  const float* campos;            /* [3]  */
  void* geom;   size_t geom_bytes;     /* >= b200gs_gauss_geom_bytes(P)       */
  void* img;    size_t img_bytes;      /* >= b200gs_gauss_image_bytes(W,H)    */
  void* binning; size_t binning_bytes; /* >= b200gs_binning_bytes(R); stage 2 */
  float* out_color;       /* [3,H,W] */
  float* out_invdepth;    /* [1,H,W] or NULL */
  float* out_normal;      /* [3,H,W] */
  float* out_refl;        /* [1,H,W] */
  int* radii;             /* [P]     */
  int exact_lists;        /* instance-list mode of this call (see b200gs_surfel_fwd) */
  int depth_slices;       /* as in b200gs_surfel_fwd: 0 automatic, 1 one slice, n <= B200GS_MAX_SLICES at most n slices */
  float slice_frac[7];    /* as in b200gs_surfel_fwd */
} b200gs_gauss_fwd;

size_t b200gs_gauss_geom_bytes(int P);
size_t b200gs_gauss_image_bytes(int width, int height);
/* The plan is the surfel path's: plan[0] = num_rendered, plan[1] = depth slices, ... (B200GS_PLAN_INTS ints). */
int b200gs_gauss_forward_stage1(const b200gs_gauss_fwd* a, int* plan, void* stream);
int b200gs_gauss_forward_stage2(const b200gs_gauss_fwd* a, const int* plan, void* stream);

/* All dL_d* outputs must be zero-filled by the caller (G/rasterize_points.cu:196-226). */
typedef struct b200gs_gauss_bwd {
  int P, sh_degree, sh_coeffs, num_rendered, width, height;
  float tan_fovx, tan_fovy, scale_modifier;
  int antialiasing, debug;
  const float* background;
  const float* means3D;
  const float* shs;               /* or NULL */
  const float* colors_precomp;    /* or NULL */
  const float* opacities;
  const float* scales;            /* or NULL */
  const float* rotations;         /* or NULL */
  const float* cov3D_precomp;     /* or NULL */
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  const int* radii;
  const void* geom; const void* binning; const void* img;
  const float* dL_dout_color;     /* [3,H,W] */
  const float* dL_dout_invdepth;  /* [1,H,W] or NULL */
  const float* dL_dout_normal;    /* [3,H,W] */
  const float* dL_dout_refl;      /* [1,H,W] or NULL (=0) */
  void* accum; size_t accum_bytes;   /* >= b200gs_gauss_accum_bytes(P) */
  float* dL_dmeans2D;     /* [P,3] "pixels" variant handed to Python (G/backward.cu:614, rasterize_points.cu:263) */
  float* dL_dconic;       /* [P,4] (x, y, -, w) or NULL */
  float* dL_dopacity;     /* [P]   */
  float* dL_dcolors;      /* [P,3] */
  float* dL_dnormals;     /* [P,3] */
  float* dL_drefl;        /* [P] or NULL */
  float* dL_dinvdepths;   /* [P] or NULL */
  float* dL_dmeans3D;     /* [P,3] */
  float* dL_dcov3D;       /* [P,6] */
  float* dL_dsh;          /* [P,M,3] or NULL */
  float* dL_dscales;      /* [P,3] */
  float* dL_drotations;   /* [P,4] */
  size_t geom_bytes, binning_bytes, img_bytes;   /* sizes of the forward's scratch as handed back (checked) */
} b200gs_gauss_bwd;

size_t b200gs_gauss_accum_bytes(int P);
int b200gs_gauss_backward(const b200gs_gauss_bwd* a, void* stream);

/* Parity helpers: GeometryState / ImageState of the 3DGS path (G/cuda_rasterizer/rasterizer_impl.h). */
int b200gs_gauss_inspect_geom(const void* geom, int P, float* depths, float* means2D /*[P,2]*/, float* cov3D /*[P,6]*/,
                              float* conic_opacity /*[P,4]*/, float* rgb /*[P,3]*/, uint32_t* tiles_touched,
                              uint32_t* point_offsets, uint8_t* clamped /*[P,3]*/, void* stream);
int b200gs_gauss_inspect_image(const void* img, int width, int height, float* accum_alpha /*[N]*/,
                               uint32_t* n_contrib /*[N]*/, uint32_t* ranges /*[T,2]*/, void* stream);

/* ------------------------------------------------------------------------ */
/* Cubemap environment lookup (C/src/cubemapencoder.h:6-17, cubemapencoder.cu:430-488,
 * :713-779).  fp32.  inputs [B,3] directions; cubemap [6,C,L,L]; fail_value [C];
 * outputs / grad_outputs are CHANNEL-MAJOR [C,B] like the reference's.  interp 0 =
 * nearest, 1 = bilinear; seamless != 0 = cross-face filtering (the mode the fork uses).
 * backward: grad_cubemap [6,C,L,L] and grad_fail [C] must be zero-filled by the caller
 * (the reference's Python does torch.zeros_like); grad_inputs [B,3] is fully written.  */
int b200gs_cubemap_forward(const float* inputs, const float* cubemap, const float* fail_value, float* outputs,
                           uint32_t interp, uint32_t seamless, uint32_t B, uint32_t C, uint32_t L, void* stream);
int b200gs_cubemap_backward(const float* grad_outputs, const float* inputs, const float* cubemap, float* grad_cubemap,
                            float* grad_inputs, float* grad_fail, uint32_t interp, uint32_t seamless, uint32_t B,
                            uint32_t C, uint32_t L, void* stream);

/* ------------------------------------------------------------------------ */
/* Per-Gaussian gradient consumer ("next" row 1): the step right after the
 * rendering path's backward.
 *
 * b200gs_adam_step replaces gaussians.optimizer.step() (train.py:306) for the groups
 * torch.optim.Adam(l, lr=0.0, eps=1e-15) is built over in scene/gaussian_model.py:196-207
 * (xyz, f_dc, f_rest, opacity, scaling, rotation, refl; the env-map group fits the same struct).
 * All groups are updated in ONE launch.  A group's gradient may be a column slice of a wider
 * row-major tensor (grad_row / grad_col): f_dc and f_rest both read the [P, M, 3] dL_dsh block
 * the backward writes, as columns [0,3) and [3,3M).  `step` is the 1-based step count shared by all
 * groups (torch keeps one per parameter; the reference steps them together).  fp32 state.  */
#define B200GS_ADAM_MAX_GROUPS 16
typedef struct {
  float* param;           /* [n] updated in place */
  const float* grad;      /* see grad_row / grad_col */
  float* exp_avg;         /* [n] */
  float* exp_avg_sq;      /* [n] */
  long long n;            /* floats in param (0: group skipped) */
  int param_row;          /* floats per row of param (0 or 1: flat) */
  int grad_row;           /* row stride of grad in floats (0: same as param_row) */
  int grad_col;           /* first column of this group inside a grad row */
  double lr;
  long long step;         /* this group's 1-based step count; 0: b200gs_adam.step (torch keeps one per parameter, and a
                           * parameter whose .grad is None in some iteration falls behind the others)          */
} b200gs_adam_group;
typedef struct {
  int groups;
  b200gs_adam_group group[B200GS_ADAM_MAX_GROUPS];
  double beta1, beta2, eps;
  long long step;         /* >= 1; used by the groups whose own step is 0 */
} b200gs_adam;
int b200gs_adam_step(const b200gs_adam* a, void* stream);

/* Densification statistics of one rendered view (train.py:242-243, scene/gaussian_model.py:579-584):
 *   where radii > 0:  max_radii2D = max(max_radii2D, radii); xyz_gradient_accum += |dL_dmeans2D|_2; denom += 1
 *   where w > 0:      accum_w += w; denom_w += 1       (skipped when gaussian_weights is NULL)
 * radii [P] int32, dL_dmeans2D [P,3], the accumulators [P] fp32 (the reference's are [P,1]).  */
int b200gs_densify_stats(int P, const int* radii, const float* dL_dmeans2D, const float* gaussian_weights,
                         float* max_radii2D, float* xyz_gradient_accum, float* denom, float* accum_w, float* denom_w,
                         void* stream);

/* ------------------------------------------------------------------------ */
/* The one collective of the path: in-place sum over the ranks of a flat fp32 buffer that
 * lives in SYMMETRIC memory (the gradient arena of a view-parallel batch), by this library's own kernel over
 * NVLink / NVSwitch peer memory.  No reference counterpart (the reference is single-GPU).
 *   multicast_ptr != NULL: the multicast address of the buffer (NVLS): rank r reduces its 1/world slice with
 *     multimem.ld_reduce and broadcasts it with multimem.st -- the adds happen in the switch;
 *   multicast_ptr == NULL: peer_ptrs (HOST array of `world` device pointers, entry i = rank i's copy as mapped in
 *     this process) are read and written directly; world in {2, 4, 8}.
 * n_floats must be a multiple of 4 and the addresses 16-B aligned.  `ctas`: thread blocks to launch (0: the library's
 * choice).  The caller orders the ranks: a device-side barrier over all ranks on `stream` before this call (all
 * arenas complete) and one after it (all slices landed). */
int b200gs_peer_allreduce_f32(void* multicast_ptr, void* const* peer_ptrs, int rank, int world, size_t n_floats,
                              int ctas, void* stream);

/* ------------------------------------------------------------------------ */
/* SSIM image loss ("next" row 3): replaces fusedssim / fusedssim_backward of
 * submodules/fused-ssim (ssim.h:5-26; called from train.py:168-169 through
 * fused_ssim/__init__.py:8-47); same map as ssim() of utils/loss_utils.py:63-101.  11x11 separable
 * Gaussian window (sigma 1.5), zero "same" padding, images [B,CH,H,W] fp32 contiguous.
 * forward: ssim_map [B,CH,H,W]; the three derivative maps (same shape) are written when all three
 * pointers are given (training) and skipped when all are NULL.
 * backward: dL_dimg1 [B,CH,H,W] from dL_dmap and the maps the forward saved.  */
int b200gs_ssim_forward(float C1, float C2, int B, int CH, int H, int W, const float* img1, const float* img2,
                        float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, void* stream);
int b200gs_ssim_backward(int B, int CH, int H, int W, const float* img1, const float* img2, const float* dL_dmap,
                         const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12, float* dL_dimg1,
                         void* stream);

/* ------------------------------------------------------------------------ */
/* Reflection shading ("next" row 2, first half): the reflection pass around the
 * rasterizer as one pixel kernel each way -- get_refl_color() + the final blend,
 * gaussian_renderer/__init__.py:22-35 and :199-201 with sample_camera_rays (utils/general_utils.py:177-197):
 *   d = normalize(R ((Kinv (x, y, 1)) - T) + R T);  refl = d - 2 n (d.n);  env = sigmoid(cubemap(refl));
 *   final = (1 - r) base + r env.
 * Kinv, R, T are HOST pointers (9, 9, 3 floats, row major; R is the camera's stored R, i.e. the array the
 * reference passes as `viewpoint_camera.R`).  normal [H,W,3]; base_color, out_final, env_color [3,H,W];
 * refl_strength [1,H,W]; cubemap [6,3,L,L]; fail_value [3]; interp / seamless as for b200gs_cubemap_forward.
 * backward: dL_denv (gradient arriving at the refl_color_map output) may be NULL; grad_cubemap
 * [6,3,L,L] and grad_fail [3] must be zero-filled by the caller; the other gradients are fully written. */
int b200gs_reflect_shade_forward(int H, int W, const float* Kinv, const float* R, const float* T, const float* normal,
                                 const float* base_color, const float* refl_strength, const float* cubemap,
                                 const float* fail_value, uint32_t L, uint32_t interp, uint32_t seamless, float* out_final,
                                 float* out_env, void* stream);
int b200gs_reflect_shade_backward(int H, int W, const float* Kinv, const float* R, const float* T, const float* normal,
                                  const float* base_color, const float* refl_strength, const float* cubemap, uint32_t L,
                                  uint32_t interp, uint32_t seamless, const float* env_color, const float* dL_dfinal,
                                  const float* dL_denv, float* dL_dbase, float* dL_drefl, float* dL_dnormal,
                                  float* grad_cubemap, float* grad_fail, void* stream);

/* Surface attributes from the rasterizer's auxiliary planes ("next" row 2, second half):
 * gaussian_renderer/__init__.py:143-183 with depths_to_points / depth_to_normal (utils/point_utils.py:9-37).
 * allmap [8,H,W] as the rasterizer returns it.  HOST constants of the camera, row major:
 *   view3x3 = world_view_transform[:3,:3] (as stored);  raymat A with rays_d = (x, y, 1) @ A
 *   (= intrins.inverse().T @ c2w[:3,:3].T of point_utils.py:20);  rays_o = c2w[:3,3].
 * forward:  out_normal [H,W,3] = rotated view normal / (norm + 1e-6);  out_surf_depth [H,W] = expected depth
 *   (allmap[0] / clamp(alpha, 1e-3)) blended with the median depth by depth_ratio, NaN/inf -> 0;
 *   out_surf_normal [3,H,W] = depth_to_normal(surf_depth) * alpha (alpha detached, border pixels zero).
 * backward: any of the three upstream gradients may be NULL; dL_dallmap [8,H,W] is fully written (planes 6, 7
 *   zero: rend_dist and the scope mask are plain slices of allmap and get their gradient from autograd). */
int b200gs_surface_forward(int H, int W, const float* allmap, const float* view3x3, const float* raymat, const float* rays_o,
                           float depth_ratio, float* out_normal, float* out_surf_depth, float* out_surf_normal, void* stream);
int b200gs_surface_backward(int H, int W, const float* allmap, const float* view3x3, const float* raymat, const float* rays_o,
                            float depth_ratio, const float* surf_depth, const float* dL_dnormal, const float* dL_dsurf_depth,
                            const float* dL_dsurf_normal, float* dL_dallmap, void* stream);

/* ------------------------------------------------------------------------ */
/* Scene IO ("next" row 4): the vertex records of point_cloud.ply <-> the per-Gaussian
 * arrays, replacing the per-property numpy walk of GaussianModel.load_ply (scene/gaussian_model.py:296-373)
 * and the concatenate + tuple list of save_ply (:239-257).  `records` is the DEVICE copy of the binary
 * vertex block: P rows of `stride` little-endian fp32 properties, exactly the bytes after "end_header\n".
 * A field is one destination array [P, cols] with `row_stride` floats between rows (row_stride > cols lets
 * features_dc / features_rest be the two column ranges of ONE packed [P,K,3] SH block, the layout the
 * preprocess kernels read); map[first + c] is the record column that holds column c of the field
 * (e.g. features_rest[:, j, ch] <- property f_rest_{ch*(K-1)+j}).  Fields must list `first` consecutively.
 * unpack: records -> fields (unmapped properties are skipped).
 * pack:   fields -> records (unmapped properties, the reference's nx ny nz, are written as 0). */
#define B200GS_PLY_MAX_FIELDS 8
#define B200GS_PLY_MAX_COLUMNS 128
typedef struct b200gs_ply_field {
  float* data;
  int cols, row_stride, first;
} b200gs_ply_field;
typedef struct b200gs_ply_layout {
  int stride, n_fields;
  b200gs_ply_field field[B200GS_PLY_MAX_FIELDS];
  unsigned char map[B200GS_PLY_MAX_COLUMNS];
} b200gs_ply_layout;
int b200gs_ply_unpack(long long P, const float* records, const b200gs_ply_layout* layout, void* stream);
int b200gs_ply_pack(long long P, const b200gs_ply_layout* layout, float* records, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200GS_H_INCLUDED */
