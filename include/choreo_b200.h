/*
 * choreo_b200.h -- C ABI of libchoreo_b200.so: the B200 (sm_100a, fp64) implementation of the
 * action-evaluation hot path of gabrielfougeron/Choreographies2 ("choreo").
 *
 * The reference has no FFI for this path: the seam is the set of numeric methods of the Cython
 * class NBodySyst (choreo/cython/_NBodySyst.pyx).  Every entry point below names the reference
 * method / cdef function it replaces (file:line in choreo/cython/_NBodySyst.pyx @ 8c8dd3f2).
 * INTEGRATION.md shows the ctypes binding a maintainer adds on the reference side.
 *
 * Conventions
 *   - all entry points are extern "C", return an int status (CHB_OK == 0) and never abort;
 *     chb_last_error_string() describes the last failure on the calling thread;
 *   - arrays are C-contiguous IEEE fp64; batched arrays carry a leading batch dimension:
 *         params   [batch][nparams]                     (reference: double[::1] params_mom_buf)
 *         segmpos  [batch][nsegm][segm_store][geodim]   (reference: double[:,:,::1] segmpos)
 *   - "_dev" operators take DEVICE pointers and are asynchronous on the given stream
 *     (a cudaStream_t passed as void*, NULL = default stream); workspace is caller-provided
 *     (chb_workspace_bytes) so one plan can serve several streams;
 *   - "_host" operators take HOST pointers, stage through plan-owned pinned/device buffers and
 *     return when the result is in the host buffer (one call at a time per plan);
 *   - a plan is immutable between chb_plan_set_nint calls; tables are copied at creation;
 *   - one plan serves ONE stream / host thread at a time: the launch bookkeeping and the caller's workspace are not duplicated per
 *     stream (several streams: one plan and one workspace each -- plans are a few hundred KB of tables).
 */
#ifndef CHOREO_B200_H
#define CHOREO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CHB_ABI_VERSION 2

enum chb_status {
    CHB_OK = 0,
    CHB_ERR_INVALID_ARGUMENT = 1, /* NULL pointer, negative size, inconsistent tables            */
    CHB_ERR_INVALID_NINT = 2,     /* nint not divisible by 2*nint_min (reference: ValueError, :545) */
    CHB_ERR_UNSUPPORTED = 3,      /* custom inter_law, geodim > 4, FFT length with a prime factor > 31 ... */
    CHB_ERR_CUDA = 4,             /* a CUDA runtime call failed (message has the CUDA error string)   */
    CHB_ERR_NO_DEVICE = 5,        /* no CUDA device / wrong architecture: there is NO CPU fallback    */
    CHB_ERR_WORKSPACE = 6,        /* workspace too small / NULL                                       */
    CHB_ERR_NOT_READY = 7         /* chb_plan_set_nint has not been called                            */
};

enum chb_inter_law {
    CHB_LAW_GRAVITY = 0, /* gravity_pot,   _NBodySyst.pyx:111-123 */
    CHB_LAW_POWER = 1    /* power_law_pot, _NBodySyst.pyx:125-133 */
};

enum chb_shortcut {
    CHB_GENERAL_SYM = 0, /* _NBodySyst.pyx:94 */
    CHB_RFFT = 1         /* _NBodySyst.pyx:95 */
};

/*
 * Flat symmetry tables = the NBodySyst fields the hot path reads (_NBodySyst.pyx:293-494, 729-737);
 * see choreographies2_b200/tables.py:extract_tables for how each is read from a live NBodySyst.
 * All pointers are HOST pointers; integer tables are int64 (Py_ssize_t) or int32 (bint / intc).
 */
typedef struct chb_tables {
    int32_t geodim;         /* NBodySyst.geodim                               */
    int32_t nbody;          /* NBodySyst.nbody                                */
    int32_t nloop;          /* NBodySyst.nloop                                */
    int32_t nsegm;          /* NBodySyst.nsegm                                */
    int32_t nbin;           /* NBodySyst.nbin_segm_unique                     */
    int32_t nint_min;       /* NBodySyst.nint_min                             */
    int32_t greater_nstore; /* NBodySyst.GreaterNStore (segm_store = segm_size + 1) */
    int32_t inter_law;      /* enum chb_inter_law                             */
    double inter_law_params[4]; /* (-nu, -nu(nu-1), nu-2, alpha), :1623-1629; gravity = (0.5,-0.75,-2.5,1) */
    /* per loop [nloop] */
    const int64_t *loopnb;     /* _loopnb        */
    const double *loopmass;    /* _loopmass      */
    const int64_t *ncoeff_min; /* _ncoeff_min_loop */
    const int64_t *n_sub_fft;  /* _n_sub_fft     */
    const int32_t *shortcut;   /* _ParamBasisShortcutPos */
    const int64_t *nnz;        /* _nnz_k_shapes[il,0]   */
    const int64_t *nppl;       /* _params_basis_shapes[il,2] */
    const int64_t *nnz_k;        /* _nnz_k_buf (flat)            */
    const int64_t *nnz_k_shifts; /* [nloop+1]                    */
    const int32_t *co_in;        /* _co_in_buf (flat)            */
    const int64_t *co_in_shifts; /* [nloop+1]                    */
    const double *params_basis;         /* _params_basis_buf_pos as interleaved (re,im), per loop C-order (geodim,nnz,nppl) */
    const int64_t *params_basis_shifts; /* [nloop+1], in complex elements */
    const int64_t *alg_iint;    /* _ALG_Iint      */
    const int64_t *alg_timerev; /* _ALG_TimeRev   */
    const double *alg_spacerot; /* _ALG_SpaceRotPos [nloop][geodim][geodim] */
    /* per body [nbody] */
    const int64_t *bodyloop; /* _bodyloop */
    /* per generating segment [nsegm] */
    const int64_t *gensegm_to_body;      /* _gensegm_to_body    */
    const int64_t *gensegm_to_iintrel;   /* _gensegm_to_iintrel */
    const int64_t *inter_timerev;        /* _InterTimeRev       */
    const double *inter_spacerot;        /* _InterSpaceRotPos [nsegm][geodim][geodim] */
    const int32_t *inter_spacerot_isid;  /* _InterSpaceRotPosIsId */
    /* per binary interaction [nbin] */
    const int64_t *bin_source;          /* _BinSourceSegm     */
    const int64_t *bin_target;          /* _BinTargetSegm     */
    const double *bin_spacerot;         /* _BinSpaceRot [nbin][geodim][geodim] */
    const int32_t *bin_spacerot_isid;   /* _BinSpaceRotIsId   */
    const double *bin_prodchargesum;    /* _BinProdChargeSum (already divided by nint_min, :818) */
    /* ABI 2, optional (all three NULL or all three given): per (body, interval of the period) the stored segment the body runs
     * on and the symmetry that carries it there -- only the post-solve reductions (chb_center_of_mass_dev, chb_detect_escape_dev)
     * read them */
    const int64_t *bodysegm;      /* _bodysegm [nbody][nint_min]                                   */
    const double *body_spacerot;  /* intersegm_to_all[ib][iint].SpaceRot [nbody][nint_min][geodim][geodim] */
    const int64_t *body_timerev;  /* intersegm_to_all[ib][iint].TimeRev  [nbody][nint_min]        */
} chb_tables;

typedef struct chb_plan chb_plan;

enum chb_query {
    CHB_Q_NINT = 0,
    CHB_Q_NPARAMS = 1,    /* NBodySyst.nparams    */
    CHB_Q_SEGM_SIZE = 2,  /* NBodySyst.segm_size  */
    CHB_Q_SEGM_STORE = 3, /* NBodySyst.segm_store */
    CHB_Q_NSEGM = 4,
    CHB_Q_GEODIM = 5,
    CHB_Q_NLOOP = 6,
    CHB_Q_NBIN = 7,
    CHB_Q_DEVICE = 8,
    CHB_Q_CHUNK = 9,          /* states processed per internal launch group             */
    CHB_Q_LAUNCHES = 10,      /* kernels launched by this plan since creation (counter)  */
    CHB_Q_FUSED = 11,         /* 0: staged kernels; else the version of the fused single-kernel path that serves this level (3 or 4) */
    CHB_Q_SEGMPOS_SIZE = 12,  /* nsegm*segm_store*geodim                                 */
    CHB_Q_STAGED_TRANSFORM = 15, /* staged path: 0 = generic per-loop kernels (any layout), 1 = parity-split kernels of planar RFFT systems */
    CHB_Q_PAIR_KERNEL = 16,      /* staged pair kernel: 0 = colour rounds (any layout), 1 = dense two-sided, 2 = tiled one-sided */
    CHB_Q_NPARAMS_INCL_O = 14, /* NBodySyst.nparams_incl_o: length of the position-parameter array of params_changevar */
    CHB_Q_HOST_POINTERS = 13  /* 1 when the *_dev entry points of this build address host memory (CPU emulation build of the
                                 test suite), 0 for the CUDA library (device pointers only)                              */
};

int chb_abi_version(void);
const char *chb_last_error_string(void);
/* number of visible CUDA devices with compute capability 10.x; 0 => every plan call fails (no CPU path) */
int chb_device_count(void);

/* NBodySyst.__init__ (tables) ; nint setter :540-590 (shapes, buffers, FFT plans) ; __dealloc__ */
int chb_plan_create(const chb_tables *tables, int device, chb_plan **plan_out);
int chb_plan_set_nint(chb_plan *plan, int64_t nint);
int chb_plan_destroy(chb_plan *plan);
int chb_plan_query(const chb_plan *plan, int what, int64_t *value_out);
/* 0 = auto; otherwise number of states per internal launch group (L2 residency knob) */
int chb_plan_set_chunk(chb_plan *plan, int64_t chunk);
/* 0 = generic staged kernels only; 4 (default) = newest fused kernel the layout supports: version 4 (rows moved by bulk
 * asynchronous copies, natural-order in-place passes) for 4096-sample RFFT loops, else version 3; 3 = force version 3
 * (A/B measurements).  Layouts versions 3 / 4 do not serve (GENERAL_SYM loops, segm_store = segm_size + 1, rotated or
 * time-reversed segments) run the general fused kernel (CHB_Q_FUSED reports 5) for any non-zero value. */
int chb_plan_set_fused(chb_plan *plan, int enable);
/* Staged transforms of planar RFFT systems (one CTA per (state, loop, residue of the time index modulo R), chb_big.cuh):
 * 0 (default) = R chosen per launch from the number of states (the smallest split that fits when the batch fills the GPU, a
 * larger one for the few live states at the end of a search); 2, 4, 8 = use that split whenever the level has it
 * (CHB_Q_STAGED_TRANSFORM reports the sum of the available R, 0 = generic kernels).  A/B measurements and tests. */
int chb_plan_set_split(chb_plan *plan, int split);

/* bytes of device workspace any *_dev operator below needs for `batch` states */
int chb_workspace_bytes(const chb_plan *plan, int64_t batch, size_t *bytes_out);

/* ---- device-pointer operators (asynchronous on `stream`) -------------------------------------- */

/* params_to_segmpos, :3607-3635 / cdef :5979-6047 */
int chb_params_to_segmpos_dev(chb_plan *plan, int64_t batch, const double *params, double *segmpos,
                              void *workspace, size_t workspace_bytes, void *stream);
/* segmpos_to_params, :3669-3696 / cdef :6121-6176 */
int chb_segmpos_to_params_dev(chb_plan *plan, int64_t batch, const double *segmpos, double *params,
                              void *workspace, size_t workspace_bytes, void *stream);
/* segmpos_to_params_T, :3699-3727 / cdef :6179-6251 */
int chb_segmpos_to_params_T_dev(chb_plan *plan, int64_t batch, const double *segmpos, double *params,
                                void *workspace, size_t workspace_bytes, void *stream);
/* params_to_kin_nrg, :2519-2528 / cdef :4918-4944 ; kin[batch] */
int chb_params_to_kin_nrg_dev(chb_plan *plan, int64_t batch, const double *params, double *kin,
                              void *workspace, size_t workspace_bytes, void *stream);
/* segm_pos_to_pot_nrg, cdef :6372-6469 ; pot[batch] */
int chb_segmpos_to_pot_nrg_dev(chb_plan *plan, int64_t batch, const double *segmpos, double *pot,
                               void *workspace, size_t workspace_bytes, void *stream);
/* segmpos_to_hash, :2542-2560 / cdef segm_pos_to_hash :6271-6369, ONE exponent per call (the reference's Hash_exp has 9):
 * hash[b] = sum over bins of BinProdChargeSum/segm_size * sum_i w_i |R x_src - x_tgt|^(2*exponent) ; hash[batch] */
int chb_segmpos_to_hash_dev(chb_plan *plan, int64_t batch, const double *segmpos, double exponent, double *hash,
                            void *workspace, size_t workspace_bytes, void *stream);
/* One Arnoldi step of the batched GMRES(restart) inside the Newton-Krylov driver, all states in one launch.  Replaces the
 * per-state inner loop of scipy.sparse.linalg.gmres that the reference reaches through choreo/scipy_plus/nonlin.py:55-134
 * (nonlin_solve_pp -> KrylovJacobian(method gmres, inner_maxiter restart 20)).  No plan state is used beyond the device.
 *   V    [>= j+2][batch][n]  Krylov bases (V[0..j] read, V[j+1] written: the new unit vector, 0 for skipped states)
 *   w    [batch][n]          Hessian-vector products A v_j of this step
 *   skip [batch] or NULL     one byte per state, non-zero = this state's GMRES has stopped
 *   h    [batch][ldh]        h[b][0..j] = column j of the Hessenberg matrix of A, h[b][j+1] = |w''|   (ldh >= j + 2)
 * The step orthogonalises w - v_j (Arnoldi on A - I; +1 is added back to h[b][j]) with modified Gram-Schmidt and a
 * second pass for the states whose vector lost more than a factor 10. */
int chb_arnoldi_step_dev(chb_plan *plan, int64_t batch, int64_t n, int j, int ldh, double *V, const double *w,
                         const unsigned char *skip, double *h, void *stream);
/* One whole inner iteration of the batched GMRES(m): chb_arnoldi_step_dev, then, per state, what the tail of an inner
 * iteration of scipy.sparse.linalg.gmres does: previous Givens rotations on the new Hessenberg column, the new rotation,
 * column j of R, the rotated right-hand side g, the iteration count and the stopping test.
 *   done   [batch] bytes     in: states to skip; out: additionally set where |g[j+1]| <= target
 *   R      [batch][m+1][m]   g [batch][m+1]   cs [batch][m][2] (c, s of rotation k)   target [batch]   kdone [batch] int64
 *   live_list_next [batch] / live_count_next [1] (both or neither NULL): the states still running after this step are
 *   appended to the list (any order; *live_count_next must be 0 on entry) for chb_..._hess_listed_dev of the next step */
int chb_gmres_step_dev(chb_plan *plan, int64_t batch, int64_t n, int j, int m, double *V, const double *w,
                       unsigned char *done, double *R, double *g, double *cs, const double *target, int64_t *kdone,
                       int *live_list_next, int *live_count_next, void *stream);
/* Norms of the Newton driver in one pass over a [batch][n] array: sumsq[b] = sum_e a[b][e]^2 (non-finite when the row has a
 * non-finite entry), maxabs[b] = max_e |a[b][e]| (may be NULL).  nonlin_solve_pp uses the 2-norm for the forcing terms and the
 * line search (nonlin.py:98,146; _safe_norm_pp) and tol_norm for the termination test (:77-78). */
int chb_row_stats_dev(chb_plan *plan, int64_t batch, int64_t n, const double *a, double *sumsq, double *maxabs, void *stream);
/* End of a GMRES cycle, all states in one launch: back substitution on each state's kdone x kdone triangle of R and
 * sol[b] = scale * sum_(j < kdone[b]) y_j V[j][b]  (scipy.sparse.linalg.gmres: solve_triangular + x += V^T y). */
int chb_gmres_solution_dev(chb_plan *plan, int64_t batch, int64_t n, int m, const double *V, const double *R,
                           const double *g, const int64_t *kdone, double scale, double *sol, void *stream);
/* segm_pos_to_pot_nrg_grad, cdef :6472-6533 ; grad_segmpos is overwritten (reference accumulates into a zeroed buffer) */
int chb_segmpos_to_pot_nrg_grad_dev(chb_plan *plan, int64_t batch, const double *segmpos, double globalmul,
                                    double *grad_segmpos, void *workspace, size_t workspace_bytes, void *stream);
/* segm_pos_to_pot_nrg_hess, cdef :6899-6962 */
int chb_segmpos_to_pot_nrg_hess_dev(chb_plan *plan, int64_t batch, const double *segmpos, const double *dsegmpos,
                                    double globalmul, double *hess_segmpos, void *workspace, size_t workspace_bytes,
                                    void *stream);
/* params_to_action, :2750-2791 ; action[batch] */
int chb_params_to_action_dev(chb_plan *plan, int64_t batch, const double *params, double *action,
                             void *workspace, size_t workspace_bytes, void *stream);
/* params_to_action_grad, :2794-2873, fused with params_to_action: action (may be NULL) and grad in one pass.
 * segmpos_out (may be NULL) receives params_to_segmpos(params) as a by-product (what
 * GetKrylovJacobian.update caches, :2013-2015). */
int chb_params_to_action_grad_dev(chb_plan *plan, int64_t batch, const double *params, double *action,
                                  double *grad, double *segmpos_out, void *workspace, size_t workspace_bytes,
                                  void *stream);
/* params_to_action_hess, :2876-2965 */
int chb_params_to_action_hess_dev(chb_plan *plan, int64_t batch, const double *params, const double *dparams,
                                  double *hess, void *workspace, size_t workspace_bytes, void *stream);
/* No counterpart in the reference.  The same Hessian-vector product for callers that
 * cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): no
 * workspace argument, and neither this call nor anything it runs allocates or frees memory.  The operator's scratch is a device
 * array built into the library, CHB_NOALLOC_SCRATCH_BYTES long (chb_noalloc_scratch_bytes reports it), that every GPU which loads
 * the library holds until exit.  Limit: the call fails with CHB_ERR_WORKSPACE, before anything is launched, when the batch needs
 * more than that (the need is what chb_workspace_bytes reports for the batch; chb_plan_set_chunk lowers it).  All calls on a
 * device share the array and find the previous call's contents in it (nothing in it has to start at zero); calls that overlap on
 * the GPU -- other streams, other plans on the same device -- race on it and must be serialised by the caller.  The first call
 * on a device looks up the array's device address (cudaGetSymbolAddress): make it once before capturing a graph. */
#define CHB_NOALLOC_SCRATCH_BYTES (32u << 20)
int chb_noalloc_scratch_bytes(size_t *bytes_out);
int chb_params_to_action_hess_noalloc_dev(chb_plan *plan, int64_t batch, const double *params, const double *dparams,
                                          double *hess, void *stream);
/* segmpos_params_to_action, :2968-2991 */
int chb_segmpos_params_to_action_dev(chb_plan *plan, int64_t batch, const double *segmpos, const double *params,
                                     double *action, void *workspace, size_t workspace_bytes, void *stream);
/* segmpos_params_to_action_grad, :2994-3041 (action may be NULL) */
int chb_segmpos_params_to_action_grad_dev(chb_plan *plan, int64_t batch, const double *segmpos,
                                          const double *params, double *action, double *grad, void *workspace,
                                          size_t workspace_bytes, void *stream);
/* segmpos_dparams_to_action_hess, :3044-3110 (the Krylov matvec, :2010-2011) */
int chb_segmpos_dparams_to_action_hess_dev(chb_plan *plan, int64_t batch, const double *segmpos,
                                           const double *dparams, double *hess, void *workspace,
                                           size_t workspace_bytes, void *stream);
/* The same operator inside the batched GMRES: skip [batch] bytes, states with a non-zero byte (their GMRES has stopped) MAY
 * be left out -- their rows of hess are then not written (the fused kernels skip them, the staged kernels compute them). */
int chb_segmpos_dparams_to_action_hess_masked_dev(chb_plan *plan, int64_t batch, const double *segmpos,
                                                  const double *dparams, const unsigned char *skip, double *hess,
                                                  void *workspace, size_t workspace_bytes, void *stream);
/* ... and with the states to process given as a list in device memory, live_list[0 .. *live_count) (each state at most
 * once, any order): the persistent CTAs share the listed states evenly.  Rows of states not listed are not written. */
int chb_segmpos_dparams_to_action_hess_listed_dev(chb_plan *plan, int64_t batch, const double *segmpos,
                                                  const double *dparams, const int *live_list, const int *live_count,
                                                  double *hess, void *workspace, size_t workspace_bytes, void *stream);

/* ---- host-pointer operators: H2D copy + operator + D2H copy, synchronous ------------------------ */
/* which: one of the CHB_OP_* codes; unused pointers may be NULL.  in0/in1 -> out0/out1:
 *   CHB_OP_PARAMS_TO_SEGMPOS       in0=params            out0=segmpos
 *   CHB_OP_SEGMPOS_TO_PARAMS       in0=segmpos           out0=params
 *   CHB_OP_SEGMPOS_TO_PARAMS_T     in0=segmpos           out0=params
 *   CHB_OP_ACTION                  in0=params            out0=action
 *   CHB_OP_ACTION_GRAD             in0=params            out0=action (or NULL) out1=grad
 *   CHB_OP_ACTION_HESS             in0=params in1=dparams out0=hess
 *   CHB_OP_SEGMPOS_ACTION          in0=segmpos in1=params out0=action
 *   CHB_OP_SEGMPOS_ACTION_GRAD     in0=segmpos in1=params out0=action (or NULL) out1=grad
 *   CHB_OP_SEGMPOS_ACTION_HESS     in0=segmpos in1=dparams out0=hess
 *   CHB_OP_KIN                     in0=params            out0=kin
 *   CHB_OP_POT                     in0=segmpos           out0=pot
 *   CHB_OP_POT_GRAD                in0=segmpos           out0=grad_segmpos   (scalar = globalmul)
 *   CHB_OP_POT_HESS                in0=segmpos in1=dsegmpos out0=hess_segmpos (scalar = globalmul)
 */
enum chb_op {
    CHB_OP_PARAMS_TO_SEGMPOS = 0,
    CHB_OP_SEGMPOS_TO_PARAMS = 1,
    CHB_OP_SEGMPOS_TO_PARAMS_T = 2,
    CHB_OP_ACTION = 3,
    CHB_OP_ACTION_GRAD = 4,
    CHB_OP_ACTION_HESS = 5,
    CHB_OP_SEGMPOS_ACTION = 6,
    CHB_OP_SEGMPOS_ACTION_GRAD = 7,
    CHB_OP_SEGMPOS_ACTION_HESS = 8,
    CHB_OP_KIN = 9,
    CHB_OP_POT = 10,
    CHB_OP_POT_GRAD = 11,
    CHB_OP_POT_HESS = 12,
    CHB_OP_HASH = 13, /* in0=segmpos out0=hash (scalar = exponent) */
    CHB_OP_PARAMS_TO_SEGMVEL = 14 /* in0=params out0=segmvel (ABI 2) */
};
int chb_run_host(chb_plan *plan, int op, int64_t batch, const double *in0, const double *in1, double scalar,
                 double *out0, double *out1);

/* ---- ODE right-hand sides of the reduced system, batched over nvec phase-space points (one time sample each) ------------
 * Compute_forces / Compute_forces_vectorized (choreo/cython/_NBodySyst.pyx:4154-4171, 4228-4246; cdef 7531-7638):
 *   pos, forces [nvec][nsegm*geodim].
 * Compute_grad_forces / Compute_grad_forces_vectorized (:4173-4191, 4249-4268; cdef 7641-7790):
 *   dpos, grad_forces [nvec][nsegm*geodim][grad_ndof].
 * Every vector is evaluated from its own positions (the per-vector definition of the reference). */
int chb_compute_forces_dev(chb_plan *plan, int64_t nvec, const double *pos, double *forces, void *stream);
int chb_compute_grad_forces_dev(chb_plan *plan, int64_t nvec, int64_t grad_ndof, const double *pos, const double *dpos,
                                double *grad_forces, void *stream);
/* host-pointer form of both (forces and / or grad_forces; dpos and grad_forces are given together or both NULL) */
int chb_compute_forces_host(chb_plan *plan, int64_t nvec, int64_t grad_ndof, const double *pos, const double *dpos,
                            double *forces, double *grad_forces);


/* ---- ABI 2: velocities, change of variables, post-solve reductions (device pointers, asynchronous on `stream`) ------------ */
/* params_to_segmvel, :3638-3666 / cdef :6050-6118 (changevar_mom_vel :4830-4914, velocity basis = i * position basis :852-854,
 * InterSpaceRotVel / ALG_SpaceRotVel = TimeRev * position rotation, _ActionSym.pyx:268-291); segmvel has the shape of segmpos */
int chb_params_to_segmvel_dev(chb_plan *plan, int64_t batch, const double *params, double *segmvel,
                              void *workspace, size_t workspace_bytes, void *stream);
/* params_changevar, :2330-2410 (changevar_mom_pos :4489-4573, _T :4746-4827, _inv :4662-4743, _invT :4576-4660).
 *   inv = 0, transpose = 0: in [batch][nparams]        -> out [batch][nparams_incl_o]
 *   inv = 0, transpose = 1: in [batch][nparams_incl_o] -> out [batch][nparams]
 *   inv = 1, transpose = 0: in [batch][nparams_incl_o] -> out [batch][nparams]
 *   inv = 1, transpose = 1: in [batch][nparams]        -> out [batch][nparams_incl_o]      (CHB_Q_NPARAMS_INCL_O) */
int chb_params_changevar_dev(chb_plan *plan, int64_t batch, const double *in, int inv, int transpose, double *out, void *stream);
/* segmpos_to_hash, :2542-2560 / cdef segm_pos_to_hash :6271-6369 with ALL exponents in one pass over the pair distances:
 * exponents is a HOST array of nexp <= 16 values (the reference's Hash_exp), hash [batch][nexp] */
int chb_segmpos_to_hash_multi_dev(chb_plan *plan, int64_t batch, const double *segmpos, int nexp, const double *exponents,
                                  double *hash, void *stream);
/* ComputeCenterOfMass, :3968-4009 ; com [batch][geodim] (needs the ABI 2 body tables) */
int chb_center_of_mass_dev(chb_plan *plan, int64_t batch, const double *segmpos, double *com, void *stream);
/* DetectEscape, :2125-2192 ; escaped [batch] int32, 1 = the graph of overlapping body boxes is not connected */
int chb_detect_escape_dev(chb_plan *plan, int64_t batch, const double *segmpos, double fac, int32_t *escaped, void *stream);
/* segm_to_path_stats, :3903-3943 / cdef :7420-7529 ; loop_len [batch][nloop], bin_dx_min [batch][nbin] */
int chb_path_stats_dev(chb_plan *plan, int64_t batch, const double *segmpos, const double *segmvel, double *loop_len,
                       double *bin_dx_min, void *stream);
/* Compute_velocities(_vectorized) / Compute_grad_velocities(_vectorized), :4081-4150 / cdef :8106-8160: out = mom / segment mass;
 * mom, out [nvec][nsegm*geodim] (grad_ndof = 0) or [nvec][nsegm*geodim][grad_ndof] */
int chb_compute_velocities_dev(chb_plan *plan, int64_t nvec, int64_t grad_ndof, const double *mom, double *out, void *stream);

/* ---- the whole batched Newton-Krylov solve inside the library ----------------------------------------------------------------
 * Replaces, for `batch` independent states at once, what the reference runs per state on the host: nonlin_solve_pp
 * (choreo/scipy_plus/nonlin.py:54-171: termination test, inexact-Newton forcing terms), its line search (:194-234 on scipy's
 * scalar_search_armijo / scalar_search_wolfe1 = MINPACK-2 dcsrch, derphi by forward difference) and scipy's
 * KrylovJacobian(gmres, restart = inner_maxiter, maxiter 1) as wired by NBodySyst.GetKrylovJacobian (_NBodySyst.pyx:2004-2032).
 * Every per-state decision is taken on the device; the host sequences launches and reads a few counters. */
typedef struct chb_newton_opts {
    double f_tol;        /* TerminationCondition(f_tol): a state is converged when norm(F) <= f_tol                          */
    double smin;         /* smallest step of the line search (nonlin_solve_pp smin; find.py linesearch_smin)                 */
    int32_t maxiter;     /* Newton iterations                                                                                */
    int32_t restart;     /* GMRES restart length = KrylovJacobian inner_maxiter (1..62)                                      */
    int32_t line_search; /* 0 none (s = smin), 1 "armijo", 2 "wolfe"                                                         */
    int32_t tol_norm;    /* 0 max norm (nonlin_solve_pp default), 1 2-norm (np.linalg.norm, what find.py passes)            */
} chb_newton_opts;
int chb_newton_workspace_bytes(const chb_plan *plan, int64_t batch, int restart, size_t *bytes_out);
/* x [batch][nparams] in: starting points, out: final iterates.  status [batch]: 1 converged, 2 iteration limit, 3 non-finite
 * residual, 4 zero Newton step (a ValueError in the reference, nonlin.py:118-121); nit [batch] Newton iterations;
 * fnorm_max / fnorm_2 [batch] (nullable) norms of the final residual; best_x [batch][nparams] / best_fnorm [batch] (nullable):
 * the iterate with the smallest residual 2-norm (current_best, nonlin.py:13-39, updated after every iteration as find.py:249-257
 * does); counters (HOST, nullable) [4]: gradient evaluations, Hessian-vector evaluations (states x calls), kernel launches,
 * Newton iterations of the batch.
 * All array pointers are device pointers; the call returns when the solve is finished. */
int chb_newton_krylov_dev(chb_plan *plan, int64_t batch, double *x, const chb_newton_opts *opts, int32_t *status, int32_t *nit,
                          double *fnorm_max, double *fnorm_2, double *best_x, double *best_fnorm, int64_t *counters,
                          void *workspace, size_t workspace_bytes, void *stream);

/* ---- memory helpers so a NumPy-only caller needs no CUDA binding ---------------------------------- */
int chb_malloc_device(chb_plan *plan, size_t bytes, void **ptr_out);
int chb_free_device(chb_plan *plan, void *ptr);
int chb_malloc_pinned(size_t bytes, void **ptr_out);
int chb_free_pinned(void *ptr);
int chb_memcpy_h2d(chb_plan *plan, void *dst_dev, const void *src_host, size_t bytes, void *stream);
int chb_memcpy_d2h(chb_plan *plan, void *dst_host, const void *src_dev, size_t bytes, void *stream);
int chb_stream_synchronize(chb_plan *plan, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* CHOREO_B200_H */
