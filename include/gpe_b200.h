/* gpe_b200 -- C ABI of the fused B200 (sm_100a) PINN train-step kernels.
 *
 * Drop-in boundary for the hot path of LevBahn/Gross-Pitaevskii-Eigenvalue-problem.  The reference
 * has no FFI of its own: its boundary is the set of nn.Module methods below, and each entry
 * point here replaces the PyTorch work those methods do (paths relative to /root/reference):
 *
 *   gpe_mlp_jet_fwd   <- nn.Sequential forward + the nested torch.autograd.grad derivatives
 *                        Gross_Pitaevskii_1D_power_Test.ipynb cell 6:50-54, 6:90-104
 *                        Gross-Pitaevskii/src/final/refine/harmonic_pinn_simulation.py:121-125, 158-172
 *                        Gross-Pitaevskii/src/gross_pitaevskii_2D.py:67-81, 183-188
 *                        Gross-Pitaevskii/src/helmholtz_2D.py:40-56, 75-83
 *   gpe_mlp_jet_bwd   <- loss.backward() through that double-backward graph
 *                        notebook cell 10:100, harmonic_pinn_simulation.py:358,
 *                        gross_pitaevskii_2D.py:376,422, helmholtz_2D.py:418
 *   gpe_gram_moments / gpe_gram_adjoints
 *                     <- pde_loss / riesz_loss / normalization_loss / symmetry_loss / loss_pde
 *                        notebook cell 6:81-162, harmonic_pinn_simulation.py:146-217,
 *                        gross_pitaevskii_1D_Different_Optimizers.py:239-375,
 *                        gross_pitaevskii_2D.py:112-213, helmholtz_2D.py:58-89
 *
 * Conventions: every pointer except the `W`/`b`/`dW`/`db` pointer ARRAYS (host arrays of device
 * pointers) and the descriptor structs is a DEVICE pointer owned by the caller; fp32, contiguous.
 * The library allocates no device memory, never synchronises the host, and enqueues all work on
 * `stream`; calls are re-entrant across streams and threads.  Gradients are WRITTEN, not accumulated.
 * Return value 0 = success, negative = error (see gpe_last_error(), thread-local).
 * There is no CPU fallback: without a CUDA device every compute entry point fails.
 *
 * Process-wide state (all of it; mutex-protected where it is written after start-up):
 *   - caches keyed by device: launch plans, SM counts, "shared-memory attribute already set" per (kernel, device);
 *   - the default of GPE_PREC_AUTO: gpe_set_tensor_path() / GPE_TENSOR (per-call control: gpe_mlp_desc.precision);
 *   - the optional profiling event list of gpe_tc_profile() / gpe_tc_timings();
 *   - tuning knobs read from the environment PER CALL: GPE_TC_TWO_TILES, GPE_TC_TWO_TILES_REV, GPE_BWD_FUSED,
 *     GPE_FUSED_NRC, GPE_RING_TILES (and once per process: GPE_TC_INTERLEAVE*, GPE_REV_PF, GPE_DW_N1, GPE_DW_PF,
 *     GPE_TC_ACT_GENERIC, GPE_NVTX).
 */
#ifndef GPE_B200_H
#define GPE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* gpe_stream_t; /* == cudaStream_t */

enum {
    GPE_ACT_TANH = 0,
    GPE_ACT_SHIFTED_TANH = 1,
    GPE_ACT_LEAKY_RELU = 2,
    /* the residual network of Gross-Pitaevskii/src/final/refine/box_to_gaussian_pinn_simulation.py:52-62 (ResidualBlock)
     * and :113-130 (build_network, use_residual=True): hidden layer 0 is Linear + ShiftedTanh; hidden layers 2k+1 and
     * 2k+2 are lin1 and lin2 of block k, out = tanh(lin2(tanh(lin1(a))) + a).  n_hidden = 1 + 2*blocks; W[] / b[] list
     * the Linear layers in module order.  d = 1, width <= 128, FP32 (FFMA) kernels. */
    GPE_ACT_RESBLOCK_TANH = 3
};
/* which derivatives travel with the value */
enum {
    GPE_JET_VALUE = 0,     /* u only                                   (boundary / mirror points)   */
    GPE_JET_DIAG = 1,      /* u, du/dx_i, d2u/dx_i^2                    h is [N,d]                   */
    GPE_JET_FULL = 2,      /* 2-D only: + d2u/dxdy                      h is [N,3] = (xx, yy, xy)    */
    GPE_JET_LAPLACIAN = 3  /* u, du/dx_i, Laplacian (one fused stream)  h is [N,1]                   */
};

/* Arithmetic of the hidden-layer contractions, chosen PER CALL by gpe_mlp_desc.precision:
 * GPE_PREC_FP32: register-blocked FFMA (fp32 throughout), any supported shape.
 * GPE_PREC_TF32X3: tcgen05 tensor cores, every fp32 operand split into hi + lo (hi = tf32) and the product formed as
 * hi*hi + (hi*lo + lo*hi), fp32 accumulation in TMEM: the hi*hi term as tf32 MMAs, the 2^-11-sized bracket as tf32 MMAs
 * (dW contraction) or as bf16 kind::f16 MMAs (forward and reverse sweeps), the forward epilogue compensating the
 * accumulator's round-toward-zero bias; fp32-accurate to ~2e-6 on the jets.  A shape without a
 * tensor-core kernel is an error (-1), never a silent change of path.
 * GPE_PREC_AUTO: TF32X3 where a tensor-core kernel exists for the shape (and the process default
 * gpe_set_tensor_path / GPE_TENSOR allows it), else FP32. */
enum { GPE_PREC_FP32 = 0, GPE_PREC_TF32X3 = 1, GPE_PREC_AUTO = 2 };

typedef struct {
    int d;        /* input dimension: 1 or 2                                     */
    int n_hidden; /* L: number of hidden layers (>= 1, <= 8)                      */
    int width;    /* W: hidden width, uniform, <= 256                             */
    int act;      /* GPE_ACT_*                                                    */
    int jet;      /* GPE_JET_*                                                    */
    int precision; /* GPE_PREC_*: arithmetic of the hidden-layer contractions      */
} gpe_mlp_desc;

/* Optional affine map applied to the inputs before the network: x_eff = a*x + b per coordinate
 * (mirror points of symmetry_loss: a=-1, b=lb+ub).  NULL = identity.  Only valid with
 * GPE_JET_VALUE. */
typedef struct {
    float a[2], b[2];
} gpe_xform;

const char* gpe_version(void);
const char* gpe_last_error(void); /* thread-local */

/* Bytes of device workspace needed by the forward / reverse call for n points. */
int gpe_workspace_bytes(const gpe_mlp_desc* desc, int64_t n, size_t* fwd_bytes, size_t* bwd_bytes);

/* W[0] [W,d], W[1..L-1] [W,W], W[L] [1,W]  (PyTorch [out,in] row-major); b[l] [W], b[L] [1].
 * x [n,d].  Outputs: u [n]; g [n,d]; h [n,H] (H by jet kind); g,h may be NULL for JET_VALUE.
 * (GPE_PREC_AUTO, GPE_JET_VALUE, n >= 16384, a shape the tensor kernels cover: the tensor forward computes the values --
 * gpe_workspace_bytes accounts for its layout.) */
int gpe_mlp_jet_fwd(const gpe_mlp_desc* desc, const float* const* W, const float* const* b, const float* x,
                    int64_t n, const gpe_xform* xform, float* u, float* g, float* h, void* ws, size_t ws_bytes,
                    gpe_stream_t stream);

/* gpe_mlp_jet_fwd without a caller workspace and without any allocation (no cudaMalloc / cudaFree / allocating
 * library call on this path).  No counterpart in the reference: added for callers
 * that cannot allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
 * The forward's scratch (packed weights + tensor-core weight images; independent of n) lives in ONE file-scope
 * `static __device__` array of GPE_STATIC_WS_BYTES (4 MiB) inside the library; every GPU whose context loads the library
 * holds that array until the process exits.  Same arguments, outputs and error codes as gpe_mlp_jet_fwd otherwise.
 * Limits: a descriptor whose gpe_workspace_bytes() forward figure exceeds GPE_STATIC_WS_BYTES, or n >
 * GPE_STATIC_MAX_POINTS, is rejected with -1 and a message (gpe_last_error).
 * Sharing: all calls on a device use the same array and it keeps the last call's contents; every byte the forward reads
 * is rewritten by this call's own pack kernel first, so nothing needs clearing between calls -- but two calls that can
 * OVERLAP on the GPU (different streams, or concurrent graph branches) race on it: serialise them on one stream.
 * The first call on a device resolves the array's address (cudaGetSymbolAddress) and sets the kernels' shared-memory
 * attributes: make that call once outside a stream capture (the usual warm-up call), later calls only launch. */
#define GPE_STATIC_WS_BYTES (4u << 20)
#define GPE_STATIC_MAX_POINTS (1ll << 24)
int gpe_mlp_jet_fwd_noalloc(const gpe_mlp_desc* desc, const float* const* W, const float* const* b, const float* x,
                            int64_t n, const gpe_xform* xform, float* u, float* g, float* h, gpe_stream_t stream);

/* Parameter gradient of  sum_n( u_bar[n] u + <g_bar[n], g> + <h_bar[n], h> ).  Any of
 * u_bar/g_bar/h_bar may be NULL (= zeros).  dW[l], db[l] have the shapes of W[l], b[l]. */
int gpe_mlp_jet_bwd(const gpe_mlp_desc* desc, const float* const* W, const float* const* b, const float* x,
                    int64_t n, const gpe_xform* xform, const float* u_bar, const float* g_bar, const float* h_bar,
                    float* const* dW, float* const* db, void* ws, size_t ws_bytes, gpe_stream_t stream);

/* ---- resident stash (memory for recompute) --------------------------------------------------
 * loss.backward() in the reference re-uses the activations autograd saved during the forward
 * (NBroot cell 10:100, S/gross_pitaevskii_2D.py:376).  gpe_mlp_jet_bwd recomputes them chunk by chunk
 * instead.  Where the caller can keep gpe_stash_bytes() of device memory alive between the forward and the
 * backward call of the SAME x, W, b (6 KB per point for [2,128x4,1]: 103 GB at 16M points, a B200 has
 * 180 GB), the forward writes the pre-activation jets once and the backward skips the recomputation.
 * *bytes = 0: this descriptor has no stash (FFMA path) -- use the plain calls. */
int gpe_stash_bytes(const gpe_mlp_desc* desc, int64_t n, size_t* bytes);
int gpe_mlp_jet_fwd_stash(const gpe_mlp_desc* desc, const float* const* W, const float* const* b, const float* x,
                          int64_t n, float* u, float* g, float* h, void* stash, size_t stash_bytes, void* ws,
                          size_t ws_bytes, gpe_stream_t stream);
int gpe_mlp_jet_bwd_stashed(const gpe_mlp_desc* desc, const float* const* W, const float* const* b, const float* x,
                            int64_t n, const float* u_bar, const float* g_bar, const float* h_bar, const void* stash,
                            size_t stash_bytes, float* const* dW, float* const* db, void* ws, size_t ws_bytes,
                            gpe_stream_t stream);

/* ---- fused loss epilogue ------------------------------------------------------------------
 * Full solution per point (1-D factor forms follow R/box_pinn_simulation.py:119-180):
 *     u   = base0 + s * F(x) * N
 *     u_i = base_i + s * (F N_i + F' N)          lap u = base_lap + s * (F lap N + 2 F' N_x + F'' N)
 * with N, N_i, lap N the network jets from gpe_mlp_jet_fwd (lap N = sum of the first `lap_cols`
 * columns of h), F = 1 or sin(pi x), `base` = caller-supplied [n, 2+d] array (value, grad, laplacian)
 * or NULL.  Features f = [1, u, lap u, V, V u, nl(u), q]; moments m[k(i,j)] = sum_n f_i f_j for
 * i <= j (row-major upper triangle, 28 entries), then m[28] = sum |grad u|^2, m[29] = sum N^2,
 * m[30] = sum (N - sym_sign * N_mirror)^2, m[31] unused.  All sums in fp64, ADDED into m
 * (the caller zeroes m; partial sums of several shards may share one buffer). */
enum { GPE_NFEAT = 7, GPE_NMOM = 32, GPE_MOM_GRAD2 = 28, GPE_MOM_NN2 = 29, GPE_MOM_SYM2 = 30 };
enum { GPE_F_ONE = 0, GPE_F_U = 1, GPE_F_LAP = 2, GPE_F_V = 3, GPE_F_VU = 4, GPE_F_NL = 5, GPE_F_Q = 6 };
enum { GPE_NL_NONE = 0, GPE_NL_POW = 1 /* u**p, integer p */, GPE_NL_ABSPOW = 2 /* |u|^(p-1) u */ };

typedef struct {
    int d;          /* 1 or 2                                                              */
    int nl_kind;    /* GPE_NL_*                                                            */
    float nl_power; /* p                                                                   */
    float scale;    /* s                                                                   */
    int bfactor;    /* 0: F = 1;  1: F = sin(pi x) (d = 1 only)                            */
    int sym_sign;   /* 0: no mirror stream; +1: (N - N_mirror)^2; -1: (N + N_mirror)^2     */
    int h_cols;     /* columns of h (0..3)                                                 */
    int lap_cols;   /* the first lap_cols columns of h sum to the Laplacian                */
} gpe_loss_desc;

/* x [n,d] (needed only with bfactor), u [n], g [n,d] or NULL, h [n,h_cols] or NULL, base [n,2+d] or
 * NULL, V [n] or NULL (=0), q [n] or NULL (=0), nm [n] (network value at the mirror point) or NULL. */
int gpe_gram_moments(const gpe_loss_desc* desc, int64_t n, const float* x, const float* u, const float* g,
                     const float* h, const float* base, const float* V, const float* q, const float* nm,
                     double* m, gpe_stream_t stream);

/* Given c[k] = dL/dm[k] (fp64, device), writes the adjoints of the NETWORK jets:
 * u_bar [n], g_bar [n,d], h_bar [n,h_cols], nm_bar [n]  (each required iff its forward input was given). */
int gpe_gram_adjoints(const gpe_loss_desc* desc, int64_t n, const float* x, const float* u, const float* g,
                      const float* h, const float* base, const float* V, const float* q, const float* nm,
                      const double* c, float* u_bar, float* g_bar, float* h_bar, float* nm_bar,
                      gpe_stream_t stream);

/* ---- fused scalar algebra of the loss terms ------------------------------------------------------
 * The loss methods of the reference end in scalar arithmetic on the reduced quantities (mu = <u,Hu>/<u,u>, mean r^2, the
 * Riesz energy, 1/(mu^2 + 1e-6) ...: NBroot cell 6:117-125, R/harmonic_pinn_simulation.py:186-196,
 * S/gross_pitaevskii_2D.py:192-211, S/helmholtz_2D.py:84-89).  Here that arithmetic is a straight-line program over the
 * moment vector(s): node i = (ops[3i], ops[3i+1], ops[3i+2]) = (opcode, a, b) with opcodes GPE_PROG_*; CONST takes
 * consts[i], INPUT reads inputs[a], the others combine earlier nodes a (and b).  Evaluated in float64 together with the
 * Jacobian d vals / d inputs (forward mode): vals [n_out], jac [n_out, n_in] row-major.  All pointers device pointers. */
enum { GPE_PROG_CONST = 0, GPE_PROG_INPUT = 1, GPE_PROG_ADD = 2, GPE_PROG_SUB = 3, GPE_PROG_MUL = 4, GPE_PROG_DIV = 5,
       GPE_PROG_NEG = 6, GPE_PROG_SQRT = 7, GPE_PROG_MAX_NODES = 1024 };
int gpe_scalar_program(const int32_t* ops, const double* consts, int n_nodes, const double* inputs, int n_in,
                       const int32_t* outs, int n_out, double* vals, double* jac, gpe_stream_t stream);

/* ---- curriculum stage bookkeeping ------------------------------------------------------------
 * What the reference's curriculum loop does on the host after every epoch
 * (Gross-Pitaevskii/src/final/refine/harmonic_pinn_simulation.py:385-423): scheduler.step(loss) of
 * CosineAnnealingWarmRestarts(T_0, T_mult, eta_min) -- the loss goes where the scheduler expects an epoch --,
 * the loss / mu / constraint logs (every epoch into rings[3][ring_width]; the reference keeps every 10th / 100th), the
 * best-loss patience counter (:394-399), the early-stop test (:401-412) and the copy of the parameters of the stop epoch.
 * All state lives in device scalars, so the epoch stays one CUDA-graph replay; one launch, one CTA.  `done` is one byte
 * (torch.bool).  The snapshot copy src[t] -> dst[t] happens only in the epoch that first trips the stop. */
#define GPE_STAGE_MAX_TENSORS 24
typedef struct {
    const float *total, *mu, *constraint; /* this epoch's scalars                       */
    float* lr;                            /* out: the next epoch's learning rate        */
    int64_t *epoch, *final_epoch, *stale; /* in/out                                     */
    unsigned char* done;                  /* in/out                                     */
    float* best;                          /* in/out                                     */
    float* rings;                         /* [3][ring_width]: total, mu, constraint     */
} gpe_stage_state;
typedef struct {
    float base_lr, tol, t_0, t_mult, eta_min;
    int64_t patience, ring_width;
} gpe_stage_consts;
typedef struct {
    int n;
    const float* src[GPE_STAGE_MAX_TENSORS];
    float* dst[GPE_STAGE_MAX_TENSORS];
    int64_t numel[GPE_STAGE_MAX_TENSORS];
} gpe_stage_snapshot;
int gpe_stage_bookkeeping(const gpe_stage_state* st, const gpe_stage_consts* c, const gpe_stage_snapshot* snap,
                          gpe_stream_t stream);

/* ---- tensor-core (tcgen05) path control ----
 * The 3xTF32 tensor path is the GPE_PREC_AUTO default wherever it applies (33 <= width <= 128 -- LeakyReLU: 65 <= width --, Laplacian jets, <= 3 hidden
 * matrices); GPE_TENSOR=0 in the environment or gpe_set_tensor_path(0) makes GPE_PREC_AUTO select the FFMA kernels instead
 * (north_star: tensor cores "only if TF32/BF16-split accuracy holds", otherwise register-blocked FFMA).
 * gpe_set_tensor_path returns the previous setting.  gpe_tc_profile(1) brackets every tensor-path backward
 * kernel with CUDA events on the launch stream; gpe_tc_timings() synchronises those events, returns the
 * accumulated milliseconds {stash forward, reverse sweep, dW contraction, fused backward} in ms[0..n) and clears the record. */
int gpe_set_tensor_path(int enable);
int gpe_tc_profile(int enable);
int gpe_tc_timings(double* ms, int n);

/* ---- FFMA peak probe (roofline denominator; MEASURED_PEAKS.json has no fp32 figure) ----
 * Runs `iters` dependent-chain-free FFMA (mode 0) or packed FFMA2 (mode 1) per thread on a
 * full grid; returns the number of fp32 FLOPs executed in *flops.  Time it with CUDA events. */
int gpe_ffma_probe(int mode, int iters, float* sink, double* flops, gpe_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* GPE_B200_H */
