// chb_api.cu -- libchoreo_b200.so: plan construction, kernel launchers and the C ABI of include/choreo_b200.h.
//
// Built by nvcc for sm_100a (product).  The same translation unit also compiles with
//   g++ -x c++ -DCHB_HOSTSIM
// into tests/hostsim/libchb_hostsim.so, where "device memory" is malloc and every kernel launch is a loop over
// blocks running the CTA programs of chb_core.cuh sequentially.  That build is test infrastructure for the
// GPU-less build container (index arithmetic / host logic); the product loader never loads it.
#include "../../include/choreo_b200.h"
#include "chb_core.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "chb_krylov.cuh"

#ifndef CHB_HOSTSIM
#include <cuda_runtime.h>
#endif
#include "chb_fused.cuh"
#include "chb_fused4.cuh"
#include "chb_fused5.cuh"
#include "chb_big.cuh"
#include "chb_ode.cuh"
#include "chb_post.cuh"
#include "chb_newton.cuh"

// ------------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_last_error;
static int chb_fail(int code, const std::string &msg) {
    g_last_error = msg;
    return code;
}
#define CHB_REQUIRE(cond, code, msg)          \
    do {                                      \
        if (!(cond)) return chb_fail(code, msg); \
    } while (0)

#ifndef CHB_HOSTSIM
#define CHB_CUDA(call)                                                                              \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess)                                                                      \
            return chb_fail(CHB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));      \
    } while (0)
typedef cudaStream_t chb_stream_t;
#else
typedef void *chb_stream_t;
#endif

// ------------------------------------------------------------------------------------------------
// plan
// ------------------------------------------------------------------------------------------------
#define CHB_HOST_SLOTS 3  // stream slots of the host-pointer pipeline (chb_run_host)
struct HostTables {
    int G = 0, nbody = 0, nloop = 0, nsegm = 0, nbin = 0, nint_min = 0, gns = 0, law = 0;
    double law_params[4] = {0, 0, 0, 0};
    std::vector<int64_t> loopnb, ncm, nsub, nnz, nppl, nnz_k, nnz_k_shifts, co_in_shifts, basis_shifts, alg_iint,
        alg_timerev, bodyloop, seg_body, seg_iintrel, seg_timerev, bin_src, bin_tgt;
    std::vector<int32_t> shortcut, co_in, seg_isid, bin_isid;
    std::vector<double> loopmass, basis, alg_rot, seg_rot, bin_rot, bin_fac;
    // ABI 2 (optional): per (body, interval) tables of the post-solve reductions
    std::vector<int64_t> bodysegm, body_timerev;
    std::vector<double> body_rot;
    int vel = 0;  // velocity plan: basis * i, rotations * TimeRev, every loop GENERAL_SYM, changevar_mom_vel scaling
};

struct chb_plan {
    int device = 0;
    HostTables T;
    int64_t nint = 0;
    bool ready = false;
    std::vector<void *> level_allocs;  // device allocations owned by the current nint level
    std::vector<ChbLoop> h_loops;      // host mirror (device pointers inside)
    ChbSys sys;                        // passed by value to kernels
    int64_t chunk_user = 0;
    int fused_enabled = 4;  // 0 staged kernels only, 1 fused v1 (separate load/store phases), 2 fused v2, 3 fused v3,
                            // 4 fused v4 where the layout allows it (chb_fused4.cuh), else v3
    int fused4_ok = 0;      // the level's layout is served by k_fused4
    int fused_resident = 148;
    int fused_active = 0;
    int64_t launches = 0;
    size_t smem_limit = 0;  // max dynamic shared memory per block (opt-in)
    int num_sms = 0;
    // per-op launch geometry
    int fft_threads = 256;
    size_t fft_smem = 0;  // dynamic smem of the transform kernels
    std::vector<int> loop_in_smem;
    int pair_threads = 256;
    size_t pair_smem[3] = {0, 0, 0};
    int pair_dense_enabled = 1;  // dense pair kernel when the level's interaction graph allows it (A/B knob: CHB_PAIR_DENSE=0)
    int dense_threads = 256;
    size_t dense_smem[3] = {0, 0, 0};
    int tile_threads = 256;
    size_t tile_smem[3] = {0, 0, 0};
    // staging for chb_run_host
    std::mutex host_mutex;
    void *stage_dev = nullptr;
    size_t stage_dev_bytes = 0;
    void *stage_pin = nullptr;
    size_t stage_pin_bytes = 0;
#ifndef CHB_HOSTSIM
    cudaStream_t host_streams[CHB_HOST_SLOTS] = {};
#endif
    ChfSys fsys;            // fused path (valid when fused_active)
    int fused_threads = 0;
    size_t fused_smem = 0;
    chb_plan *vel_plan = nullptr;  // sibling plan of params_to_segmvel (same nint level), owned by this plan
    ChbBodyTab body;               // device copies of the body tables (valid when body.nbody > 0)
    const long long *d_pos_off = nullptr;  // [nloop + 1] offsets of the loops in position-parameter space (params_changevar)
    long long nparams_incl_o = 0;
    int big_active = 0;     // staged transforms by the parity-split kernels of chb_big.cuh (planar RFFT systems without a fused kernel)
    ChbBigSys bigsys[3];    // residue splits R = 2, 4, 8 (index log2 R - 1); big_ok[i] says which exist at this level
    size_t big_smem[3] = {0, 0, 0};
    int big_ok[3] = {0, 0, 0};
    int big_min_len[3] = {0, 0, 0};  // shortest transform length of each split
    const unsigned char *stage_skip = nullptr;  // set by the Newton driver around staged launches: byte mask of states to leave out
    int64_t stage_c0 = 0;                       // first state of the chunk being launched (offset into stage_skip)
    int gmres_split_hint = 1;  // 0: the Newton driver has few live states -> Givens tail inside the Arnoldi launch (one launch less per inner iteration)
    int big_force_r = 0;    // CHB_BIG_R: A/B knob, use this split whenever it exists
    int fused5_active = 0;  // the level is served by the general fused kernel (chb_fused5.cuh) instead of versions 3 / 4
    Ch5Sys f5;              // passed by value (constant bank) to k_fused5
};

// ---- memory shims --------------------------------------------------------------------------------
static int dev_alloc(chb_plan *p, size_t bytes, void **out) {
    if (bytes == 0) bytes = 16;
#ifndef CHB_HOSTSIM
    CHB_CUDA(cudaSetDevice(p->device));
    CHB_CUDA(cudaMalloc(out, bytes));
#else
    (void)p;
    *out = malloc(bytes);
    CHB_REQUIRE(*out, CHB_ERR_CUDA, "malloc failed");
#endif
    return CHB_OK;
}
static void dev_free(void *ptr) {
#ifndef CHB_HOSTSIM
    cudaFree(ptr);
#else
    free(ptr);
#endif
}
static int dev_upload(void *dst, const void *src, size_t bytes) {
    if (bytes == 0) return CHB_OK;
#ifndef CHB_HOSTSIM
    CHB_CUDA(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
#else
    memcpy(dst, src, bytes);
#endif
    return CHB_OK;
}
template <class T>
static int level_upload(chb_plan *p, const std::vector<T> &v, const T **out) {
    void *d = nullptr;
    int rc = dev_alloc(p, v.size() * sizeof(T), &d);
    if (rc) return rc;
    p->level_allocs.push_back(d);
    rc = dev_upload(d, v.data(), v.size() * sizeof(T));
    if (rc) return rc;
    *out = (const T *)d;
    return CHB_OK;
}
static void free_level(chb_plan *p) {
    for (void *q : p->level_allocs) dev_free(q);
    p->level_allocs.clear();
    p->h_loops.clear();
    p->ready = false;
}

// ------------------------------------------------------------------------------------------------
// C ABI: misc
// ------------------------------------------------------------------------------------------------
extern "C" int chb_abi_version(void) { return CHB_ABI_VERSION; }
extern "C" const char *chb_last_error_string(void) { return g_last_error.c_str(); }

extern "C" int chb_device_count(void) {
#ifndef CHB_HOSTSIM
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    int ok = 0;
    for (int i = 0; i < n; ++i) {
        cudaDeviceProp pr;
        if (cudaGetDeviceProperties(&pr, i) == cudaSuccess && pr.major == 10) ++ok;
    }
    return ok;
#else
    return 1;
#endif
}

template <class T>
static void copy_in(std::vector<T> &dst, const T *src, size_t n) {
    dst.assign(src, src + n);
}

extern "C" int chb_plan_create(const chb_tables *t, int device, chb_plan **out) {
    CHB_REQUIRE(t && out, CHB_ERR_INVALID_ARGUMENT, "chb_plan_create: NULL argument");
    CHB_REQUIRE(t->geodim >= 1 && t->geodim <= CHB_MAX_G, CHB_ERR_UNSUPPORTED,
                "chb_plan_create: geodim must be in [1,4]");
    CHB_REQUIRE(t->nloop >= 1 && t->nsegm >= 1 && t->nbin >= 0 && t->nint_min >= 1 && t->nbody >= 1,
                CHB_ERR_INVALID_ARGUMENT, "chb_plan_create: bad sizes");
    CHB_REQUIRE(t->inter_law == CHB_LAW_GRAVITY || t->inter_law == CHB_LAW_POWER, CHB_ERR_UNSUPPORTED,
                "chb_plan_create: only gravity_pot and power_law_pot are supported on the device");
#ifndef CHB_HOSTSIM
    {
        int n = 0;
        cudaError_t e = cudaGetDeviceCount(&n);
        if (e != cudaSuccess || n == 0) {
            cudaGetLastError();
            return chb_fail(CHB_ERR_NO_DEVICE, "chb_plan_create: no CUDA device visible (there is no CPU fallback)");
        }
        CHB_REQUIRE(device >= 0 && device < n, CHB_ERR_INVALID_ARGUMENT, "chb_plan_create: bad device index");
        cudaDeviceProp pr;
        CHB_CUDA(cudaGetDeviceProperties(&pr, device));
        CHB_REQUIRE(pr.major == 10, CHB_ERR_NO_DEVICE,
                    "chb_plan_create: device is not compute capability 10.x (library is built for sm_100a only)");
    }
#endif
    chb_plan *p = new chb_plan();
    p->device = device;
    HostTables &T = p->T;
    T.G = t->geodim;
    T.nbody = t->nbody;
    T.nloop = t->nloop;
    T.nsegm = t->nsegm;
    T.nbin = t->nbin;
    T.nint_min = t->nint_min;
    T.gns = t->greater_nstore ? 1 : 0;
    T.law = t->inter_law;
    for (int i = 0; i < 4; ++i) T.law_params[i] = t->inter_law_params[i];
    const int nl = T.nloop, G = T.G;
    copy_in(T.loopnb, t->loopnb, nl);
    copy_in(T.loopmass, t->loopmass, nl);
    copy_in(T.ncm, t->ncoeff_min, nl);
    copy_in(T.nsub, t->n_sub_fft, nl);
    copy_in(T.shortcut, t->shortcut, nl);
    copy_in(T.nnz, t->nnz, nl);
    copy_in(T.nppl, t->nppl, nl);
    copy_in(T.nnz_k_shifts, t->nnz_k_shifts, nl + 1);
    copy_in(T.co_in_shifts, t->co_in_shifts, nl + 1);
    copy_in(T.basis_shifts, t->params_basis_shifts, nl + 1);
    copy_in(T.nnz_k, t->nnz_k, (size_t)T.nnz_k_shifts[nl]);
    copy_in(T.co_in, t->co_in, (size_t)T.co_in_shifts[nl]);
    copy_in(T.basis, t->params_basis, (size_t)(2 * T.basis_shifts[nl]));
    copy_in(T.alg_iint, t->alg_iint, nl);
    copy_in(T.alg_timerev, t->alg_timerev, nl);
    copy_in(T.alg_rot, t->alg_spacerot, (size_t)nl * G * G);
    copy_in(T.bodyloop, t->bodyloop, T.nbody);
    copy_in(T.seg_body, t->gensegm_to_body, T.nsegm);
    copy_in(T.seg_iintrel, t->gensegm_to_iintrel, T.nsegm);
    copy_in(T.seg_timerev, t->inter_timerev, T.nsegm);
    copy_in(T.seg_rot, t->inter_spacerot, (size_t)T.nsegm * G * G);
    copy_in(T.seg_isid, t->inter_spacerot_isid, T.nsegm);
    copy_in(T.bin_src, t->bin_source, T.nbin);
    copy_in(T.bin_tgt, t->bin_target, T.nbin);
    copy_in(T.bin_rot, t->bin_spacerot, (size_t)T.nbin * G * G);
    copy_in(T.bin_isid, t->bin_spacerot_isid, T.nbin);
    copy_in(T.bin_fac, t->bin_prodchargesum, T.nbin);
    if (t->bodysegm && t->body_spacerot && t->body_timerev) {
        const size_t nbi = (size_t)T.nbody * T.nint_min;
        copy_in(T.bodysegm, t->bodysegm, nbi);
        copy_in(T.body_timerev, t->body_timerev, nbi);
        copy_in(T.body_rot, t->body_spacerot, nbi * G * G);
    }
    // consistency checks (the reference would hit undefined behaviour instead, asserts compiled out: setup.py:237-239)
    std::string err;
    for (int il = 0; il < nl && err.empty(); ++il) {
        int64_t nnz = T.nnz[il], nppl = T.nppl[il];
        if (nnz < 0 || nppl < 0 || nnz > CHB_MAX_NNZ || nppl > CHB_MAX_NPPL || nnz * nppl > CHB_MAX_C)
            err = "loop with nnz*nppl beyond compiled limits";
        else if (T.nnz_k_shifts[il + 1] - T.nnz_k_shifts[il] != nnz)
            err = "nnz_k_shifts inconsistent with nnz";
        else if (T.basis_shifts[il + 1] - T.basis_shifts[il] != (int64_t)G * nnz * nppl)
            err = "params_basis_shifts inconsistent";
        else if (T.nsub[il] != 1 && T.nsub[il] != 2)
            err = "Impossible value for n_sub_fft (allowed: 1 or 2)";
        else if (T.ncm[il] < 1 || T.loopnb[il] < 1 || !(T.loopmass[il] != 0.0))
            err = "bad ncoeff_min / loopnb / loopmass";
        else if (T.shortcut[il] == CHB_RFFT) {
            bool ok = nnz == 1 && nppl == 2 * G && T.nnz_k[T.nnz_k_shifts[il]] == 0 && T.nsub[il] == 1;
            if (!ok) err = "RFFT shortcut requires nnz=1, nnz_k=[0], nppl=2*geodim, n_sub_fft=1";
        } else if (T.shortcut[il] != CHB_GENERAL_SYM)
            err = "unknown shortcut";
        if (err.empty() && nnz > 0 && T.nnz_k[T.nnz_k_shifts[il]] == 0) {
            if (T.co_in_shifts[il + 1] - T.co_in_shifts[il] != nppl) err = "co_in must have nppl entries when nnz_k[0]==0";
        }
    }
    for (int s = 0; s < T.nsegm && err.empty(); ++s) {
        if (T.seg_body[s] < 0 || T.seg_body[s] >= T.nbody) err = "gensegm_to_body out of range";
        else if (T.bodyloop[T.seg_body[s]] < 0 || T.bodyloop[T.seg_body[s]] >= nl) err = "bodyloop out of range";
        else if (T.seg_timerev[s] != 1 && T.seg_timerev[s] != -1) err = "InterTimeRev must be +-1";
        else if (T.seg_iintrel[s] < 0) err = "gensegm_to_iintrel negative";
    }
    for (int b = 0; b < T.nbin && err.empty(); ++b) {
        if (T.bin_src[b] < 0 || T.bin_src[b] >= T.nsegm || T.bin_tgt[b] < 0 || T.bin_tgt[b] >= T.nsegm)
            err = "BinSourceSegm/BinTargetSegm out of range";
    }
    for (size_t k = 0; k < T.bodysegm.size() && err.empty(); ++k) {
        if (T.bodysegm[k] < 0 || T.bodysegm[k] >= T.nsegm) err = "bodysegm out of range";
        else if (T.body_timerev[k] != 1 && T.body_timerev[k] != -1) err = "body TimeRev must be +-1";
    }
    if (!err.empty()) {
        delete p;
        return chb_fail(CHB_ERR_INVALID_ARGUMENT, "chb_plan_create: " + err);
    }
#ifndef CHB_HOSTSIM
    {
        cudaDeviceProp pr;
        cudaGetDeviceProperties(&pr, device);
        p->smem_limit = pr.sharedMemPerBlockOptin;
        p->num_sms = pr.multiProcessorCount;
    }
#else
    p->smem_limit = 227 * 1024;
    p->num_sms = 148;
#endif
    memset(&p->sys, 0, sizeof(p->sys));
    *out = p;
    return CHB_OK;
}

extern "C" int chb_plan_destroy(chb_plan *p) {
    if (!p) return CHB_OK;
    if (p->vel_plan) chb_plan_destroy(p->vel_plan);
    p->vel_plan = nullptr;
    free_level(p);
    if (p->stage_dev) dev_free(p->stage_dev);
#ifndef CHB_HOSTSIM
    for (int i = 0; i < CHB_HOST_SLOTS; ++i)
        if (p->host_streams[i]) cudaStreamDestroy(p->host_streams[i]);
    if (p->stage_pin) cudaFreeHost(p->stage_pin);
#else
    if (p->stage_pin) free(p->stage_pin);
#endif
    delete p;
    return CHB_OK;
}

// ---- FFT planning -----------------------------------------------------------------------------------
static bool factorize(int M, std::vector<int> &fac) {
    fac.clear();
    int m = M;
    while (m % 4 == 0) { fac.push_back(4); m /= 4; }
    while (m % 2 == 0) { fac.push_back(2); m /= 2; }
    for (int p = 3; p <= CHB_MAX_PRIME && m > 1; p += 2)
        while (m % p == 0) { fac.push_back(p); m /= p; }
    if (m != 1) return false;
    if ((int)fac.size() > CHB_MAX_FAC) return false;
    return true;
}

static void exact_root(long long num, long long den, double *c, double *s) {
    // exp(-2 pi i num/den): exact integer octant reduction, then long double cos/sin on [0, pi/4]
    num %= den;
    if (num < 0) num += den;
    const long double tau = 6.283185307179586476925286766559005768L;
    const long long a = 8 * num;  // theta = 2 pi a / (8 den)
    const int oct = (int)(a / den);
    long long phi;
    switch (oct) {
        default:
        case 0: phi = a; break;
        case 1: phi = 2 * den - a; break;
        case 2: phi = a - 2 * den; break;
        case 3: phi = 4 * den - a; break;
        case 4: phi = a - 4 * den; break;
        case 5: phi = 6 * den - a; break;
        case 6: phi = a - 6 * den; break;
        case 7: phi = 8 * den - a; break;
    }
    const long double ang = tau * (long double)phi / (long double)(8 * den);
    const long double cp = cosl(ang), sp = sinl(ang);
    long double cr, sr;
    switch (oct) {
        default:
        case 0: cr = cp; sr = sp; break;
        case 1: cr = sp; sr = cp; break;
        case 2: cr = -sp; sr = cp; break;
        case 3: cr = -cp; sr = sp; break;
        case 4: cr = -cp; sr = -sp; break;
        case 5: cr = -sp; sr = -cp; break;
        case 6: cr = sp; sr = -cp; break;
        case 7: cr = cp; sr = -sp; break;
    }
    *c = (double)cr;
    *s = (double)(-sr);  // exp(-i theta)
}

struct FftCache {
    const cd *roots;
    const int *rev;
    std::vector<int> fac;
};


// ---- fused path planning (chb_fused.cuh) ------------------------------------------------------------
static bool chf_factorize(int N, std::vector<int> &fac) {
    fac.clear();
    // odd radices first (outermost), the largest power of two last: the last factor is the pass that the v2 kernel
    // fuses with the parameter load / result store (chf_load_pass / chf_store_pass), which needs it even
    int m = N, e2 = 0;
    while (m % 2 == 0) { m /= 2; ++e2; }
    for (int q : {7, 5, 3})
        while (m % q == 0) { fac.push_back(q); m /= q; }
    const int rem = e2 % 4;
    if (rem == 1) fac.push_back(2);
    if (rem == 2) fac.push_back(4);
    if (rem == 3) fac.push_back(8);
    for (; e2 >= 4; e2 -= 4) fac.push_back(16);
    return m == 1 && !fac.empty() && (int)fac.size() <= CHF_MAX_FAC;
}

// returns 1 when the fused path can serve this level (and fills p->fsys), 0 when not, <0 on error
static int fused_setup(chb_plan *p, const std::vector<ChbLoop> &loops, int segm_size) {
    const HostTables &T = p->T;
    static_assert(sizeof(ChfSys) + sizeof(ChfArgs) <= 4000, "kernel parameter space");
    if (T.G != 2 || T.gns || T.nloop > CHF_MAX_LOOPS || T.nsegm > CHF_MAX_SEGS || T.nsegm < 1) return 0;
    ChfSys &F = p->fsys;
    memset(&F, 0, sizeof(F));
    F.nloop = T.nloop;
    F.nseg = T.nsegm;
    F.segm_size = segm_size;
    F.law_kind = T.law;
    for (int i = 0; i < 4; ++i) F.law[i] = T.law_params[i];
    F.nparams = p->sys.nparams;
    F.segsz = p->sys.segsz;
    int aligned = (F.nparams % 2 == 0) ? 1 : 0;
    int smem_cd = 0, maxN = 0;
    std::map<int, int> first_loop_of_N;
    for (int il = 0; il < T.nloop; ++il) {
        const ChbLoop &L = loops[il];
        if (L.shortcut != CHB_RFFT || L.nnz != 1 || L.nppl != 4 || L.n_sub_fft != 1 || !L.has_k0) return 0;
        ChfLoop &FL = F.loops[il];
        FL.N = 2 * L.npr;
        FL.npr = L.npr;
        if (FL.N % segm_size != 0) return 0;
        std::vector<int> fac;
        if (!chf_factorize(FL.N, fac)) return 0;
        FL.nfac = (int)fac.size();
        for (int i = 0; i < FL.nfac; ++i) FL.fac[i] = fac[i];
        FL.swm = (FL.N % 16 == 0) ? 15 : 0;
        FL.smem_off = smem_cd;
        smem_cd += FL.N;
        maxN = std::max(maxN, FL.N);
        FL.mom_off = L.mom_off;
        FL.nco_in = L.nco_in;
        FL.ncm = L.ncm;
        FL.co0 = L.co_in[0] ? L.co_prefix[0] : -1;
        FL.co2 = L.co_in[2] ? L.co_prefix[2] : -1;
        FL.loopmul = L.loopmul;
        if ((L.mom_off - L.nco_in) % 2 != 0) aligned = 0;
        // every slot of the loop must be covered by exactly one generating segment
        std::vector<int> cover(FL.N / segm_size, 0);
        for (int s = 0; s < T.nsegm; ++s)
            if ((int)T.bodyloop[T.seg_body[s]] == il) {
                if (T.seg_iintrel[s] >= (int64_t)cover.size()) return 0;
                cover[T.seg_iintrel[s]]++;
            }
        for (int c : cover)
            if (c != 1) return 0;
    }
    F.aligned16 = aligned;
    for (int s = 0; s < T.nsegm; ++s) {
        ChfSeg &sg = F.segs[s];
        sg.loop = (int)T.bodyloop[T.seg_body[s]];
        sg.row0 = (int)T.seg_iintrel[s] * segm_size;
        sg.timerev = (int)T.seg_timerev[s];
        sg.isid = T.seg_isid[s] ? 1 : 0;
        for (int i = 0; i < 4; ++i) sg.rot[i] = T.seg_rot[(size_t)s * 4 + i];
    }
    for (int b = 0; b < T.nbin; ++b) {
        ChfPair &P = F.pairs[T.bin_src[b] * CHF_MAX_SEGS + T.bin_tgt[b]];
        if (P.on) return 0;  // two interactions on the same ordered segment pair: generic path
        P.on = 1;
        P.isid = T.bin_isid[b] ? 1 : 0;
        P.fac = T.bin_fac[b];
        for (int i = 0; i < 4; ++i) P.rot[i] = T.bin_rot[(size_t)b * 4 + i];
    }
    {   // common layout served by chf_pair_fast
        bool fast = T.law == 0;
        for (int s = 0; s < T.nsegm && fast; ++s) {
            const ChfSeg &sg = F.segs[s];
            const ChfLoop &FL = F.loops[sg.loop];
            fast = sg.isid && sg.timerev > 0;
            (void)FL;
            for (int t = 0; t < T.nsegm && fast; ++t) {
                const ChfPair &P = F.pairs[s * CHF_MAX_SEGS + t];
                fast = t > s ? (P.on && P.isid) : !P.on;
            }
        }
        F.pair_fast = fast ? 1 : 0;
    }
    // threads: one group per loop, about one radix-16 butterfly pair per thread and pass
    int Tg = ((maxN / 32 + 31) / 32) * 32;
    int cap = (CHF_MAX_THREADS / T.nloop) / 32 * 32;
    Tg = std::max(32, std::min(Tg, cap));
    if (cap < 32) return 0;
    F.T = Tg;
    p->fused_threads = Tg * T.nloop;
    p->fused_smem = CHF_RED_BYTES + (size_t)smem_cd * sizeof(cd);
    if (p->fused_smem > p->smem_limit) return 0;
    {   // co-resident CTAs = distance (in states) of the L2 prefetch
        const size_t per_cta = p->fused_smem + 1024;
        int occ = (int)std::min<size_t>((size_t)(228 * 1024) / per_cta, (size_t)(2048 / p->fused_threads));
        occ = std::max(1, std::min(occ, 32));
        F.pf_stride = occ * p->num_sms;
        p->fused_resident = occ * p->num_sms;
        // tensor-memory stash of ver 3: 8 words per parameter row, R/2 rows per butterfly, one butterfly per pair iteration
        int cols_thread = 0;
        for (int il = 0; il < T.nloop; ++il) {
            const ChfLoop &FL = F.loops[il];
            const int R = FL.fac[FL.nfac - 1], M = FL.N / R, npairs = (M + 1) / 2, half = Tg / 2;
            cols_thread = std::max(cols_thread, ((npairs + half - 1) / half) * 4 * R);
        }
        const int blocks = (p->fused_threads / 32 + 3) / 4;
        int cols = 32;
        while (cols < blocks * cols_thread) cols *= 2;
        F.tmem_thread_cols = cols_thread;
        F.tmem_cols = (cols <= 512 && occ * cols <= 512 && cols_thread <= 512) ? cols : 0;
    }
    // tables, shared between loops of equal length
    for (int il = 0; il < T.nloop; ++il) {
        ChfLoop &FL = F.loops[il];
        auto it = first_loop_of_N.find(FL.N);
        if (it != first_loop_of_N.end()) {
            FL.rev = F.loops[it->second].rev;
            for (int i = 0; i < FL.nfac; ++i) FL.tw[i] = F.loops[it->second].tw[i];
            continue;
        }
        first_loop_of_N[FL.N] = il;
        const int N = FL.N;
        std::vector<int> rev(N);
        for (int k = 0; k < N; ++k) {
            int kk = k, pos = 0, Lcur = N;
            for (int i = 0; i < FL.nfac; ++i) {
                int r = FL.fac[i];
                int t = kk % r;
                kk /= r;
                Lcur /= r;
                pos += t * Lcur;
            }
            rev[k] = pos;
        }
        int rc = level_upload(p, rev, &FL.rev);
        if (rc) return -1;
        int Lcur = N;
        for (int i = 0; i < FL.nfac; ++i) {
            const int r = FL.fac[i], Lp = Lcur / r, nb = chf_nbase(r);
            std::vector<cd> tw((size_t)std::max(1, Lp * nb));
            for (int j = 0; j < Lp; ++j)
                for (int e = 0; e < nb; ++e) {
                    double c, sn;
                    exact_root((long long)j * (1LL << e), Lcur, &c, &sn);
                    tw[(size_t)e * Lp + j] = mk(c, sn);  // [e][j]: conflict-free / coalesced for consecutive j
                }
            rc = level_upload(p, tw, &FL.tw[i]);
            if (rc) return -1;
            FL.tw_len[i] = Lp > 1 ? Lp * nb : 0;
            FL.tw_soff[i] = -1;
            Lcur = Lp;
        }
    }
    for (int il = 0; il < T.nloop; ++il) {
        ChfLoop &FL = F.loops[il];
        const int src = first_loop_of_N[FL.N];
        for (int i = 0; i < FL.nfac; ++i) { FL.tw_len[i] = F.loops[src].tw_len[i]; FL.tw_soff[i] = -1; }
    }
    {   // shared-memory copy of the twiddle tables of the longest loop's length when they fit behind the loop buffers
        int src = 0;
        for (int il = 1; il < T.nloop; ++il)
            if (F.loops[il].N > F.loops[src].N) src = il;
        src = first_loop_of_N[F.loops[src].N];
        int need = 0;
        for (int i = 0; i < F.loops[src].nfac; ++i) need += F.loops[src].tw_len[i];
        const size_t total = p->fused_smem + (size_t)need * sizeof(cd);
        F.tw_smem_len = 0;
        if (need > 0 && total <= p->smem_limit && (size_t)(228 * 1024) / (total + 1024) >= (size_t)std::max(1, p->fused_resident / p->num_sms)) {
            int off = 0;
            for (int il = 0; il < T.nloop; ++il) {
                if (F.loops[il].N != F.loops[src].N) continue;
                int o = 0;
                for (int i = 0; i < F.loops[il].nfac; ++i) {
                    F.loops[il].tw_soff[i] = F.loops[il].tw_len[i] > 0 ? o : -1;
                    o += F.loops[il].tw_len[i];
                }
                off = o;
            }
            F.tw_src_loop = src;
            F.tw_smem_off = smem_cd;
            F.tw_smem_len = off;
            p->fused_smem = total;
        }
    }
    {   // version 4 (chb_fused4.cuh): every loop 4096 = 16^3 samples covered by one identity segment, gravity fast pairs,
        // 16-byte aligned rows (bulk copies), tables in shared memory, the tensor-memory stash available
        bool ok = F.pair_fast && F.aligned16 && F.tmem_cols > 0 && F.tw_smem_len > 0 && T.nsegm == T.nloop && T.nloop <= 3 &&
                  Tg == CH4_T && segm_size == CH4_N;
        for (int il = 0; il < T.nloop && ok; ++il) {
            const ChfLoop &FL = F.loops[il];
            ok = FL.N == CH4_N && FL.nfac == 3 && FL.fac[0] == 16 && FL.fac[1] == 16 && FL.fac[2] == 16 && FL.nco_in % 2 == 0 &&
                 FL.nco_in <= 2 && FL.tw_soff[0] >= 0 && FL.tw_soff[1] >= 0;
        }
        for (int s = 0; s < T.nsegm && ok; ++s) ok = F.segs[s].row0 == 0 && F.segs[s].loop == s;
        if (ok && p->fused_smem + 256 * sizeof(cd) + CHF_RED_BYTES + 512 <= p->smem_limit) {
            // full 16 x 16 twiddle table of pass "k1", then the second buffer of partial sums
            F.tw4_off = (int)((p->fused_smem - CHF_RED_BYTES) / sizeof(cd));
            p->fused_smem += 256 * sizeof(cd) + CHF_RED_BYTES;
        } else {
            ok = false;
        }
        p->fused4_ok = ok ? 1 : 0;
    }
    return 1;
}


// factors of N for the general fused kernel, LARGEST radix first: the table of a DIF stage has (L / r) * nbase(r) entries, so
// the long first stages get the big radices and the tables of all stages fit in shared memory up to N of about 10^4
static bool ch5_factorize(int N, std::vector<int> &fac) {
    fac.clear();
    int m = N, e2 = 0;
    while (m % 2 == 0) { m /= 2; ++e2; }
    for (; e2 >= 4; e2 -= 4) fac.push_back(16);
    if (e2 == 3) fac.push_back(8);
    if (e2 == 2) fac.push_back(4);
    if (e2 == 1) fac.push_back(2);
    for (int q : {7, 5, 3})
        while (m % q == 0) { fac.push_back(q); m /= q; }
    return m == 1 && !fac.empty() && (int)fac.size() <= CHF_MAX_FAC;
}

// ---- general fused path planning (chb_fused5.cuh) ----------------------------------------------------
// returns 1 when k_fused5 can serve this level (and fills p->f5), 0 when not, <0 on error
static int fused5_setup(chb_plan *p, const std::vector<ChbLoop> &loops, int segm_size, int segm_store) {
    const HostTables &T = p->T;
    static_assert(sizeof(Ch5Sys) + sizeof(ChfArgs) <= 32000, "kernel parameter space (CUDA >= 12.1: 32764 bytes)");
    if (T.G != 2 || T.nloop > CHF_MAX_LOOPS || T.nsegm > CH5_MAX_SEGS || T.nsegm < 1) return 0;
    Ch5Sys &F = p->f5;
    memset(&F, 0, sizeof(F));
    F.nloop = T.nloop;
    F.nseg = T.nsegm;
    F.segm_size = segm_size;
    F.segm_store = segm_store;
    F.law_kind = T.law;
    for (int i = 0; i < 4; ++i) F.law[i] = T.law_params[i];
    F.nparams = p->sys.nparams;
    F.segsz = p->sys.segsz;
    long long smem_cd = 0;
    int maxN = 0;
    std::map<int, int> first_loop_of_N;
    for (int il = 0; il < T.nloop; ++il) {
        const ChbLoop &L = loops[il];
        Ch5Loop &FL = F.loops[il];
        if (L.nnz < 1 || L.nnz > CH5_MAX_NNZ || L.nppl > CHB_MAX_NPPL) return 0;
        FL.kind = L.shortcut;
        FL.n_sub_fft = L.n_sub_fft;
        FL.nnz = L.nnz;
        FL.nppl = L.nppl;
        FL.C = L.C;
        FL.has_k0 = L.has_k0;
        if (L.shortcut == CHB_RFFT) {
            if (L.nnz != 1 || L.nppl != 4 || L.n_sub_fft != 1 || !L.has_k0) return 0;
            FL.Cp = 1;
        } else {
            if (L.C < 2 || (L.C & 1) || L.C > CH5_MAX_C) return 0;
            FL.Cp = L.C / 2;
        }
        ChfLoop &f = FL.f;
        f.N = 2 * L.npr;
        f.npr = L.npr;
        std::vector<int> fac;
        if (!ch5_factorize(f.N, fac)) return 0;
        f.nfac = (int)fac.size();
        for (int i = 0; i < f.nfac; ++i) f.fac[i] = fac[i];
        f.swm = (f.N % 16 == 0) ? 15 : 0;
        f.mom_off = L.mom_off;
        f.nco_in = L.nco_in;
        f.ncm = L.ncm;
        f.co0 = (L.shortcut == CHB_RFFT && L.co_in[0]) ? L.co_prefix[0] : -1;
        f.co2 = (L.shortcut == CHB_RFFT && L.co_in[2]) ? L.co_prefix[2] : -1;
        f.loopmul = L.loopmul;
        FL.seq_stride = f.N;
        f.smem_off = (int)smem_cd;
        smem_cd += (long long)FL.Cp * f.N + 16;  // + the Adjust_after_last_gen row, keeps the next loop 256-byte aligned
        maxN = std::max(maxN, f.N);
        for (int j = 0; j < L.nppl; ++j) { FL.co_in[j] = L.co_in[j]; FL.co_prefix[j] = L.co_prefix[j]; }
        for (int c = 0; c < L.C; ++c) FL.ikc[c] = c / L.nppl;
        FL.tw_trivial = 1;
        for (int ik = 0; ik < L.nnz; ++ik) {
            FL.nnz_k[ik] = L.nnz_k[ik];
            FL.wmir[ik] = L.wmir[ik];
            if (L.nnz_k[ik] != 0) FL.tw_trivial = 0;
        }
        for (int g = 0; g < 2; ++g)
            for (int c = 0; c < L.C; ++c) { FL.bre[g][c] = L.bre[g][c]; FL.bim[g][c] = L.bim[g][c]; }
        FL.gen_tw = L.gen_tw;
        FL.has_alg = L.has_alg;
        FL.alg_src = L.alg_src;
        for (int i = 0; i < 4; ++i) FL.alg_rot[i] = L.alg_rot[i];
        // rows produced by the transform and their coverage: every slot of segm_size rows holds exactly one segment
        const int main_rows = (L.n_sub_fft == 2) ? L.npr : 2 * L.npr;
        FL.nrows = (L.n_sub_fft == 2) ? L.npr + 1 : 2 * L.npr;
        if (main_rows % segm_size != 0) return 0;
        if (segm_store != segm_size && L.n_sub_fft == 1 && !L.has_alg) return 0;
        std::vector<int> cover(main_rows / segm_size, 0);
        for (int s = 0; s < T.nsegm; ++s)
            if ((int)T.bodyloop[T.seg_body[s]] == il) {
                if (T.seg_iintrel[s] >= (int64_t)cover.size()) return 0;
                cover[T.seg_iintrel[s]]++;
            }
        for (int c : cover)
            if (c != 1) return 0;
    }
    for (int s = 0; s < T.nsegm; ++s) {
        ChfSeg &sg = F.segs[s];
        sg.loop = (int)T.bodyloop[T.seg_body[s]];
        sg.row0 = (int)T.seg_iintrel[s] * segm_size;
        sg.timerev = (int)T.seg_timerev[s];
        sg.isid = T.seg_isid[s] ? 1 : 0;
        for (int i = 0; i < 4; ++i) sg.rot[i] = T.seg_rot[(size_t)s * 4 + i];
    }
    for (int b = 0; b < T.nbin; ++b) {
        ChfPair &P = F.pairs[T.bin_src[b] * CH5_MAX_SEGS + T.bin_tgt[b]];
        if (P.on) return 0;  // two interactions on the same ordered segment pair: generic path
        P.on = 1;
        P.isid = T.bin_isid[b] ? 1 : 0;
        P.fac = T.bin_fac[b];
        for (int i = 0; i < 4; ++i) P.rot[i] = T.bin_rot[(size_t)b * 4 + i];
    }
    int Tg = ((maxN / 32 + 31) / 32) * 32;
    const int cap = (CHF_MAX_THREADS / T.nloop) / 32 * 32;
    if (cap < 32) return 0;
    Tg = std::max(32, std::min(Tg, cap));
    F.T = Tg;
    p->fused_threads = Tg * T.nloop;
    p->fused_smem = CHF_RED_BYTES + (size_t)smem_cd * sizeof(cd);
    if (p->fused_smem > p->smem_limit) return 0;
    {
        const size_t per_cta = p->fused_smem + 1024;
        int occ = (int)std::min<size_t>((size_t)(228 * 1024) / per_cta, (size_t)(2048 / p->fused_threads));
        occ = std::max(1, std::min(occ, 32));
        p->fused_resident = occ * p->num_sms;
    }
    // FFT tables (global memory), shared between loops of equal length
    for (int il = 0; il < T.nloop; ++il) {
        ChfLoop &f = F.loops[il].f;
        auto it = first_loop_of_N.find(f.N);
        if (it != first_loop_of_N.end()) {
            const ChfLoop &g = F.loops[it->second].f;
            f.rev = g.rev;
            for (int i = 0; i < f.nfac; ++i) { f.tw[i] = g.tw[i]; f.tw_len[i] = g.tw_len[i]; f.tw_soff[i] = -1; }
            continue;
        }
        first_loop_of_N[f.N] = il;
        const int N = f.N;
        std::vector<int> rev(N);
        for (int k = 0; k < N; ++k) {
            int kk = k, pos = 0, Lcur = N;
            for (int i = 0; i < f.nfac; ++i) {
                const int r = f.fac[i], t = kk % r;
                kk /= r;
                Lcur /= r;
                pos += t * Lcur;
            }
            rev[k] = pos;
        }
        if (level_upload(p, rev, &f.rev)) return -1;
        int Lcur = N;
        for (int i = 0; i < f.nfac; ++i) {
            const int r = f.fac[i], Lp = Lcur / r, nb = chf_nbase(r);
            std::vector<cd> tw((size_t)std::max(1, Lp * nb));
            for (int j = 0; j < Lp; ++j)
                for (int e = 0; e < nb; ++e) {
                    double c, sn;
                    exact_root((long long)j * (1LL << e), Lcur, &c, &sn);
                    tw[(size_t)e * Lp + j] = mk(c, sn);
                }
            if (level_upload(p, tw, &f.tw[i])) return -1;
            f.tw_len[i] = Lp > 1 ? Lp * nb : 0;
            f.tw_soff[i] = -1;
            Lcur = Lp;
        }
    }
    {   // shared-memory copies of the stage tables, smallest first, while they fit behind the loop buffers
        struct Item { int il, s, len; };
        std::vector<Item> items;
        for (auto &kv : first_loop_of_N) {
            const ChfLoop &f = F.loops[kv.second].f;
            for (int i = 0; i < f.nfac; ++i)
                if (f.tw_len[i] > 0) items.push_back({kv.second, i, f.tw_len[i]});
        }
        std::sort(items.begin(), items.end(), [](const Item &a, const Item &b) { return a.len < b.len; });
        size_t room = p->smem_limit > p->fused_smem + 1024 ? (p->smem_limit - p->fused_smem - 1024) / sizeof(cd) : 0;
        int off = 0;
        for (const Item &it : items) {
            if ((size_t)it.len > room) break;
            F.loops[it.il].f.tw_soff[it.s] = off;
            off += it.len;
            room -= it.len;
        }
        for (int il = 0; il < T.nloop; ++il) {  // loops of equal length share the tables
            const int src = first_loop_of_N[F.loops[il].f.N];
            for (int i = 0; i < F.loops[il].f.nfac; ++i) F.loops[il].f.tw_soff[i] = F.loops[src].f.tw_soff[i];
        }
        F.tw_smem_off = (int)smem_cd;
        F.tw_smem_len = off;
        p->fused_smem += (size_t)off * sizeof(cd);
    }
    return 1;
}

// ---- parity-split staged transforms (chb_big.cuh) ----------------------------------------------------
// returns 1 when k_big_fwd<R> / k_big_adj<R> can serve the level's transforms (and fills F, *smem), 0 when not, <0 on error
static int big_setup(chb_plan *p, const std::vector<ChbLoop> &loops, int segm_size, int segm_store, int R, ChbBigSys &F, size_t *smem) {
    const HostTables &T = p->T;
    if (T.G != 2 || segm_store != segm_size || T.nloop > CHB_BIG_MAX_LOOPS || T.vel) return 0;
    memset(&F, 0, sizeof(F));
    F.R = R;
    F.nloop = T.nloop;
    F.segm_size = segm_size;
    F.nparams = p->sys.nparams;
    F.segsz = p->sys.segsz;
    int maxNh = 0;
    std::map<int, int> first_loop_of_N;
    for (int il = 0; il < T.nloop; ++il) {
        const ChbLoop &L = loops[il];
        if (L.shortcut != CHB_RFFT || L.nnz != 1 || L.nppl != 4 || L.n_sub_fft != 1 || !L.has_k0) return 0;
        const int N = 2 * L.npr, Nh = N / R;
        if (N % R != 0 || Nh < 2 || N % segm_size != 0) return 0;
        ChfLoop &h = F.loops[il].h;
        h.N = Nh;
        h.npr = L.npr;
        std::vector<int> fac;
        if (!ch5_factorize(Nh, fac)) return 0;
        h.nfac = (int)fac.size();
        for (int i = 0; i < h.nfac; ++i) h.fac[i] = fac[i];
        h.swm = (Nh % 16 == 0) ? 15 : 0;
        h.mom_off = L.mom_off;
        h.nco_in = L.nco_in;
        h.ncm = L.ncm;
        h.co0 = L.co_in[0] ? L.co_prefix[0] : -1;
        h.co2 = L.co_in[2] ? L.co_prefix[2] : -1;
        h.loopmul = L.loopmul;
        maxNh = std::max(maxNh, Nh);
        std::vector<int> slot_seg(N / segm_size, -1);
        for (int s = 0; s < T.nsegm; ++s)
            if ((int)T.bodyloop[T.seg_body[s]] == il) {
                if (T.seg_iintrel[s] >= (int64_t)slot_seg.size() || slot_seg[T.seg_iintrel[s]] >= 0) return 0;
                slot_seg[T.seg_iintrel[s]] = s;
            }
        for (int v : slot_seg)
            if (v < 0) return 0;
        if (level_upload(p, slot_seg, &F.loops[il].slot_seg)) return -1;
    }
    // one radix-16 butterfly per thread and pass when the half length allows it: these kernels are latency-bound (one CTA per SM
    // because of the buffer), so the warps per SM set their speed; CHB_BIG_T caps the block size (A/B knob, read once)
    static const int big_t_cap = [] { const char *e = getenv("CHB_BIG_T"); const int v = e ? atoi(e) : 0; return v >= 64 ? std::min(v, CHB_BIG_MAX_THREADS) : CHB_BIG_MAX_THREADS; }();
    int Tg = ((maxNh / 16 + 31) / 32) * 32;
    Tg = std::max(64, std::min(Tg, big_t_cap));
    F.T = Tg;
    int max_slots = 0;
    for (int il = 0; il < T.nloop; ++il) max_slots = std::max(max_slots, (R * F.loops[il].h.N) / segm_size);
    F.max_slots = max_slots;
    static const int stagger_ns = [] { const char *e = getenv("CHB_BIG_STAGGER_NS"); return e ? atoi(e) : 0; }();  // A/B knob, read once
    F.stagger_ns = stagger_ns;
    F.first_wave = p->num_sms;
    const size_t slot_bytes = (size_t)max_slots * sizeof(ChbBigSlot);
    size_t tw_room = 0;
    {
        const size_t base = CHF_RED_BYTES + (size_t)maxNh * sizeof(cd) + slot_bytes;
        if (base + 1024 > p->smem_limit) return 0;
        tw_room = (p->smem_limit - base - 1024) / sizeof(cd);
    }
    int max_tw = 0;
    for (int il = 0; il < T.nloop; ++il) {
        ChbBigLoop &BL = F.loops[il];
        ChfLoop &h = BL.h;
        auto it = first_loop_of_N.find(h.N);
        if (it != first_loop_of_N.end()) {
            const ChbBigLoop &G0 = F.loops[it->second];
            h.rev = G0.h.rev;
            for (int i = 0; i < h.nfac; ++i) { h.tw[i] = G0.h.tw[i]; h.tw_len[i] = G0.h.tw_len[i]; h.tw_soff[i] = G0.h.tw_soff[i]; }
            BL.t_hi = G0.t_hi;
            BL.t_lo = G0.t_lo;
            continue;
        }
        first_loop_of_N[h.N] = il;
        const int Nh = h.N, N = R * Nh;
        std::vector<int> rev(Nh);
        for (int k = 0; k < Nh; ++k) {
            int kk = k, pos = 0, Lcur = Nh;
            for (int i = 0; i < h.nfac; ++i) {
                const int r = h.fac[i], t = kk % r;
                kk /= r;
                Lcur /= r;
                pos += t * Lcur;
            }
            rev[k] = pos;
        }
        if (level_upload(p, rev, &h.rev)) return -1;
        int Lcur = Nh, off = 0;
        size_t room = tw_room;
        for (int i = 0; i < h.nfac; ++i) {
            const int r = h.fac[i], Lp = Lcur / r, nb = chf_nbase(r);
            std::vector<cd> tw((size_t)std::max(1, Lp * nb));
            for (int j = 0; j < Lp; ++j)
                for (int e = 0; e < nb; ++e) {
                    double c, sn;
                    exact_root((long long)j * (1LL << e), Lcur, &c, &sn);
                    tw[(size_t)e * Lp + j] = mk(c, sn);
                }
            if (level_upload(p, tw, &h.tw[i])) return -1;
            h.tw_len[i] = Lp > 1 ? Lp * nb : 0;
            h.tw_soff[i] = -1;
            if (h.tw_len[i] > 0 && (size_t)h.tw_len[i] <= room) {  // stage table also in shared memory
                h.tw_soff[i] = off;
                off += h.tw_len[i];
                room -= h.tw_len[i];
            }
            Lcur = Lp;
        }
        max_tw = std::max(max_tw, off);
        std::vector<cd> thi((size_t)(N + 127) / 128 + 1), tlo(128);
        for (size_t j = 0; j < thi.size(); ++j) {
            double c, sn;
            exact_root((long long)j * 128, N, &c, &sn);
            thi[j] = mk(c, sn);
        }
        for (int j = 0; j < 128; ++j) {
            double c, sn;
            exact_root(j, N, &c, &sn);
            tlo[j] = mk(c, sn);
        }
        if (level_upload(p, thi, &BL.t_hi) || level_upload(p, tlo, &BL.t_lo)) return -1;
    }
    F.tw_smem_len = max_tw;
    *smem = CHF_RED_BYTES + ((size_t)maxNh + max_tw) * sizeof(cd) + slot_bytes;
    return *smem <= p->smem_limit ? 1 : 0;
}

extern "C" int chb_plan_set_nint(chb_plan *p, int64_t nint) {
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, "chb_plan_set_nint: NULL plan");
    HostTables &T = p->T;
    if (nint <= 0 || nint % (2 * (int64_t)T.nint_min) != 0)
        return chb_fail(CHB_ERR_INVALID_NINT, "Provided nint " + std::to_string(nint) + " should be divisible by " +
                                                  std::to_string(2 * T.nint_min));
    CHB_REQUIRE(nint < (1LL << 30), CHB_ERR_UNSUPPORTED, "chb_plan_set_nint: nint too large");
    for (int il = 0; il < T.nloop; ++il)
        CHB_REQUIRE(nint % (2 * T.ncm[il]) == 0, CHB_ERR_INVALID_NINT, "nint not divisible by 2*ncoeff_min_loop");
    free_level(p);
    const int G = T.G, nl = T.nloop;
    const int segm_size = (int)(nint / T.nint_min);
    const int segm_store = segm_size + (T.gns ? 1 : 0);

    std::map<int, FftCache> fft_cache;
    std::vector<ChbLoop> loops(nl);
    long long mom_off = 0, scratch_off = 0;
    std::vector<int> seg_loop(T.nsegm);
    for (int s = 0; s < T.nsegm; ++s) seg_loop[s] = (int)T.bodyloop[T.seg_body[s]];

    for (int il = 0; il < nl; ++il) {
        ChbLoop &L = loops[il];
        memset(&L, 0, sizeof(L));
        L.shortcut = T.shortcut[il];
        L.n_sub_fft = (int)T.nsub[il];
        L.nnz = (int)T.nnz[il];
        L.nppl = (int)T.nppl[il];
        L.C = L.nnz * L.nppl;
        L.ncm = (int)T.ncm[il];
        L.npr = (int)(nint / (2 * T.ncm[il]));
        L.ncomp = (L.n_sub_fft == 2) ? L.npr + 1 : 2 * L.npr;
        L.ninter = L.ncomp + ((L.n_sub_fft == 1 && T.gns) ? 1 : 0);
        L.has_k0 = (L.nnz > 0 && T.nnz_k[T.nnz_k_shifts[il]] == 0) ? 1 : 0;
        L.nseq = (L.shortcut == CHB_RFFT) ? G : L.C;
        L.stride = L.npr + 1;
        L.mom_off = mom_off;
        L.loopmul = 1.0 / sqrt((double)T.loopnb[il] * T.loopmass[il] * 4.0 * M_PI * M_PI);
        L.vel = T.vel;
        L.velmul = 1.0 / sqrt((double)T.loopnb[il] * T.loopmass[il]);
        L.invmass = 1.0 / T.loopmass[il];
        int ncor = 0;
        for (int j = 0; j < L.nppl; ++j) {
            int keep = 1;
            if (L.has_k0) keep = T.co_in[T.co_in_shifts[il] + j] ? 1 : 0;
            L.co_in[j] = keep;
            L.co_prefix[j] = ncor;
            ncor += keep;
        }
        L.nco_in = L.has_k0 ? (L.nppl - ncor) : 0;
        mom_off += (long long)L.npr * L.C - L.nco_in;
        for (int ik = 0; ik < L.nnz; ++ik) {
            L.nnz_k[ik] = (int)T.nnz_k[T.nnz_k_shifts[il] + ik];
            double c, s;
            exact_root(L.nnz_k[ik], L.ncm, &c, &s);
            L.wmir[ik] = mk(c, -s);  // exp(+2 pi i nnz_k/ncm)
        }
        const double *bas = T.basis.data() + 2 * T.basis_shifts[il];
        for (int g = 0; g < G; ++g)
            for (int c = 0; c < L.C; ++c) {
                L.bre[g][c] = bas[2 * ((long long)g * L.C + c)];
                L.bim[g][c] = bas[2 * ((long long)g * L.C + c) + 1];
            }
        L.has_alg = (segm_size != segm_store && L.nnz > 0 && L.n_sub_fft == 1) ? 1 : 0;
        L.alg_src = T.alg_timerev[il] == 1 ? (int)(T.alg_iint[il] * segm_size)
                                            : (int)((T.alg_iint[il] + 1) * segm_size - 1);
        if (L.has_alg && (L.alg_src < 0 || L.alg_src >= L.ncomp)) {
            free_level(p);
            return chb_fail(CHB_ERR_INVALID_ARGUMENT, "chb_plan_set_nint: ALG source row out of range");
        }
        for (int i = 0; i < G * G; ++i) L.alg_rot[i] = T.alg_rot[(size_t)il * G * G + i];
        L.scratch_off = scratch_off;
        L.smem_cd = (long long)L.nseq * L.stride;
        scratch_off += L.smem_cd;

        if (L.nnz > 0) {
            const int M = L.npr;
            auto it = fft_cache.find(M);
            if (it == fft_cache.end()) {
                FftCache fc;
                if (!factorize(M, fc.fac)) {
                    free_level(p);
                    return chb_fail(CHB_ERR_UNSUPPORTED, "chb_plan_set_nint: FFT length " + std::to_string(2 * M) +
                                                             " has a prime factor > " + std::to_string(CHB_MAX_PRIME));
                }
                const int N = 2 * M;
                std::vector<cd> roots(N);
                for (int j = 0; j < N; ++j) {
                    double c, s;
                    exact_root(j, N, &c, &s);
                    roots[j] = mk(c, s);
                }
                std::vector<int> rev(M);
                for (int k = 0; k < M; ++k) {
                    int kk = k, pos = 0, Lcur = M;
                    for (int r : fc.fac) {
                        int t = kk % r;
                        kk /= r;
                        Lcur /= r;
                        pos += t * Lcur;
                    }
                    rev[k] = pos;
                }
                int rc = level_upload(p, roots, &fc.roots);
                if (!rc) rc = level_upload(p, rev, &fc.rev);
                if (rc) { free_level(p); return rc; }
                it = fft_cache.insert(std::make_pair(M, fc)).first;
            }
            L.fft.M = M;
            L.fft.N = 2 * M;
            L.fft.nfac = (int)it->second.fac.size();
            for (int i = 0; i < L.fft.nfac; ++i) L.fft.fac[i] = it->second.fac[i];
            L.fft.roots = it->second.roots;
            L.fft.rev = it->second.rev;
            if (L.shortcut == CHB_GENERAL_SYM) {
                std::vector<cd> tw((size_t)(L.npr + 1) * L.nnz);
                for (int m = 0; m <= L.npr; ++m)
                    for (int ik = 0; ik < L.nnz; ++ik) {
                        double c, s;
                        exact_root((long long)m * L.nnz_k[ik], nint, &c, &s);
                        tw[(size_t)m * L.nnz + ik] = mk(c, s);
                    }
                int rc = level_upload(p, tw, &L.gen_tw);
                if (rc) { free_level(p); return rc; }
            }
        }
        // segments of this loop + CSR over slice slots
        std::vector<int> segs;
        int max_slot = -1;
        for (int s = 0; s < T.nsegm; ++s)
            if (seg_loop[s] == il) {
                segs.push_back(s);
                max_slot = std::max(max_slot, (int)T.seg_iintrel[s]);
                long long last_row = T.seg_iintrel[s] * (long long)segm_size + segm_store - 1;
                if (last_row >= L.ninter) {
                    free_level(p);
                    return chb_fail(CHB_ERR_INVALID_ARGUMENT, "chb_plan_set_nint: segment reads past its loop slice");
                }
            }
        L.seg_count = (int)segs.size();
        L.n_iint = max_slot + 1;
        std::vector<int> off(L.n_iint + 1, 0), lst;
        for (int q = 0; q < L.n_iint; ++q) {
            for (int s : segs)
                if (T.seg_iintrel[s] == q) lst.push_back(s);
            off[q + 1] = (int)lst.size();
        }
        int rc = level_upload(p, segs, &L.segs);
        if (!rc) rc = level_upload(p, off, &L.iint_off);
        if (!rc) rc = level_upload(p, lst, &L.iint_seg);
        if (rc) { free_level(p); return rc; }
    }

    // segments / bins
    std::vector<ChbSeg> segs(T.nsegm);
    for (int s = 0; s < T.nsegm; ++s) {
        memset(&segs[s], 0, sizeof(ChbSeg));
        segs[s].loop = seg_loop[s];
        segs[s].iintrel = (int)T.seg_iintrel[s];
        segs[s].timerev = (int)T.seg_timerev[s];
        segs[s].isid = T.seg_isid[s] ? 1 : 0;
        for (int i = 0; i < G * G; ++i) segs[s].rot[i] = T.seg_rot[(size_t)s * G * G + i];
    }
    // greedy edge colouring of the interaction multigraph: a round never touches a segment twice
    std::vector<int> colour(T.nbin, -1);
    int nround = 0;
    {
        std::vector<std::vector<char>> used;  // used[round][segment]
        for (int b = 0; b < T.nbin; ++b) {
            int s = (int)T.bin_src[b], t = (int)T.bin_tgt[b];
            int r = 0;
            for (;; ++r) {
                if (r == (int)used.size()) used.push_back(std::vector<char>(T.nsegm, 0));
                if (!used[r][s] && !used[r][t]) break;
            }
            used[r][s] = used[r][t] = 1;
            colour[b] = r;
        }
        nround = (int)used.size();
    }
    std::vector<ChbBin> bins;
    std::vector<int> round_off(nround + 1, 0);
    int max_round_bins = 1;
    for (int r = 0; r < nround; ++r) {
        for (int b = 0; b < T.nbin; ++b)
            if (colour[b] == r) {
                ChbBin B;
                memset(&B, 0, sizeof(B));
                B.src = (int)T.bin_src[b];
                B.tgt = (int)T.bin_tgt[b];
                B.isid = T.bin_isid[b] ? 1 : 0;
                B.fac = T.bin_fac[b];
                B.pad = b;  // position of the interaction in the caller's order (chb_path_stats_dev)
                for (int i = 0; i < G * G; ++i) B.rot[i] = T.bin_rot[(size_t)b * G * G + i];
                bins.push_back(B);
            }
        round_off[r + 1] = (int)bins.size();
        max_round_bins = std::max(max_round_bins, round_off[r + 1] - round_off[r]);
    }

    ChbSys &S = p->sys;
    memset(&S, 0, sizeof(S));
    S.G = G;
    S.nloop = nl;
    S.nsegm = T.nsegm;
    S.nbin = T.nbin;
    S.segm_size = segm_size;
    S.segm_store = segm_store;
    S.nparams = mom_off;
    S.segsz = (long long)T.nsegm * segm_store * G;
    S.law_kind = T.law;
    for (int i = 0; i < 4; ++i) S.law[i] = T.law_params[i];
    S.nround = nround;
    S.scratch_per_state = scratch_off;
    int rc = level_upload(p, loops, &S.loops);
    if (!rc) rc = level_upload(p, segs, &S.segs);
    if (!rc) rc = level_upload(p, bins, &S.bins);
    if (!rc) rc = level_upload(p, round_off, &S.round_off);
    if (rc) { free_level(p); return rc; }
    p->h_loops = loops;
    {   // position-parameter offsets (params_changevar) and the body tables of the post-solve reductions
        std::vector<long long> pos_off(nl + 1, 0);
        for (int il = 0; il < nl; ++il) pos_off[il + 1] = pos_off[il] + (long long)loops[il].npr * loops[il].C;
        p->nparams_incl_o = 2 * pos_off[nl];
        rc = level_upload(p, pos_off, &p->d_pos_off);
        memset(&p->body, 0, sizeof(p->body));
        if (!rc && !T.bodysegm.empty()) {
            const size_t nbi = (size_t)T.nbody * T.nint_min;
            std::vector<int> bs(nbi), tr(nbi);
            std::vector<double> mass(T.nbody);
            double tot = 0.0;
            for (size_t k = 0; k < nbi; ++k) { bs[k] = (int)T.bodysegm[k]; tr[k] = (int)T.body_timerev[k]; }
            for (int ib = 0; ib < T.nbody; ++ib) {
                mass[ib] = T.loopmass[T.bodyloop[ib]];
                tot += mass[ib] * T.nint_min;
            }
            rc = level_upload(p, bs, &p->body.bodysegm);
            if (!rc) rc = level_upload(p, tr, &p->body.timerev);
            if (!rc) rc = level_upload(p, T.body_rot, &p->body.rot);
            if (!rc) rc = level_upload(p, mass, &p->body.bodymass);
            p->body.nbody = T.nbody;
            p->body.nint_min = T.nint_min;
            p->body.totmass = tot;
        }
        if (rc) { free_level(p); return rc; }
    }

    // ---- launch geometry ----
    // transform kernels: FFT buffer in shared memory when it fits, else in the per-state global scratch
    p->fft_threads = 256;
    const size_t red_bytes = sizeof(double) * 1024;
    size_t budget = p->smem_limit > red_bytes + 1024 ? p->smem_limit - red_bytes - 1024 : 0;
    p->loop_in_smem.assign(nl, 0);
    size_t need = 0;
    long long maxM = 1;
    for (int il = 0; il < nl; ++il) {
        size_t b = (size_t)loops[il].smem_cd * sizeof(cd);
        if (b <= budget) {
            p->loop_in_smem[il] = 1;
            need = std::max(need, b);
        }
        maxM = std::max<long long>(maxM, (long long)loops[il].nseq * loops[il].npr);
    }
    if (maxM >= 4096) p->fft_threads = 512;
    p->fft_smem = red_bytes + need;
    // pair kernel tile: shared arrays spos (+sdpos) + sout, thread item = (bin in round, sample)
    {
        size_t per_sample = (size_t)T.nsegm * G * sizeof(double);
        size_t target = 64 * 1024;  // for 3 arrays
        int tile = 256;
        while (tile > 1 && per_sample * tile * 3 > target) tile >>= 1;
        while (tile > 1 && tile / 2 >= segm_store) tile >>= 1;
        if (per_sample * tile * 3 + red_bytes > p->smem_limit) {
            free_level(p);
            return chb_fail(CHB_ERR_UNSUPPORTED, "chb_plan_set_nint: too many segments for the pair kernel tile");
        }
        S.pair_tile = tile;
        S.pair_ntiles = (segm_store + tile - 1) / tile;
        long long items = (long long)max_round_bins * tile;
        int thr = 64;
        while (thr < 1024 && thr < items) thr <<= 1;
        p->pair_threads = thr;
        size_t arr = per_sample * tile;
        p->pair_smem[0] = red_bytes + arr;       // energy only: spos
        p->pair_smem[1] = red_bytes + 2 * arr;   // + sout
        p->pair_smem[2] = red_bytes + 3 * arr;   // + sdpos
        // dense pair kernel (cta_pair_dense): planar, identity BinSpaceRot everywhere, no self-interaction, and enough
        // interactions per segment that evaluating every pair from both sides beats the coloured scatter
        bool dense = G == 2 && T.nsegm >= 2 && T.nbin >= 2 * T.nsegm;
        for (int b = 0; b < T.nbin && dense; ++b) dense = T.bin_isid[b] && T.bin_src[b] != T.bin_tgt[b];
        const int spad = (T.nsegm + 31) / 32 * 32;
        if (dense && spad <= 256) {
            std::vector<double> Wm((size_t)T.nsegm * spad, 0.0);
            for (int b = 0; b < T.nbin; ++b) {
                Wm[(size_t)T.bin_src[b] * spad + T.bin_tgt[b]] += T.bin_fac[b];
                Wm[(size_t)T.bin_tgt[b] * spad + T.bin_src[b]] += T.bin_fac[b];
            }
            int rcw = level_upload(p, Wm, &S.pair_w);
            if (rcw) { free_level(p); return rcw; }
            S.pair_spad = spad;
            {   // every pair of distinct segments interacts with the same weight?
                bool uni = true;
                const double w0 = Wm[1];
                for (int a = 0; a < T.nsegm && uni; ++a)
                    for (int b = 0; b < T.nsegm && uni; ++b)
                        if (a != b) uni = Wm[(size_t)a * spad + b] == w0;
                S.pair_uniform = (uni && w0 != 0.0 && !(getenv("CHB_PAIR_UNIFORM") && atoi(getenv("CHB_PAIR_UNIFORM")) == 0)) ? 1 : 0;
                S.pair_wu = w0;
            }
            const size_t arrd = (size_t)tile * spad * sizeof(cd), wbytes = (size_t)T.nsegm * spad * sizeof(double);
            p->dense_threads = 256 / spad * spad >= spad ? 256 : spad;
            p->dense_smem[0] = red_bytes + wbytes + arrd;
            p->dense_smem[1] = red_bytes + wbytes + 2 * arrd;
            p->dense_smem[2] = red_bytes + wbytes + 3 * arrd;
            S.pair_dense = p->dense_smem[2] <= p->smem_limit ? 1 : 0;
        }
        if (getenv("CHB_PAIR_DENSE") && atoi(getenv("CHB_PAIR_DENSE")) == 0) S.pair_dense = 0;
        // tiled one-sided kernel (cta_pair_tile) on the same conditions, up to 64 segments: every unordered pair once, 17 FP64
        // instructions per unordered pair instead of 30.  Its first version lost to the dense kernel (22 % FP64 pipe utilisation:
        // the run-time law branch made every pair evaluation a basic block of its own and pulled pow() into the unrolled tile);
        // with the law as a template parameter it wins (profiles/r02_pair_tile_law0_ab.txt: 64C 0.592 -> 0.493 ms per 256 states,
        // 16C 0.465 -> 0.358 ms per 1024 states) and is the default; CHB_PAIR_TILE=0 selects the dense kernel (A/B).
        S.pair_tilek = 0;
        static const int tile_on = [] { const char *e = getenv("CHB_PAIR_TILE"); return e ? atoi(e) : 1; }();
        if (S.pair_dense && T.nsegm <= 64 && tile_on == 1) {
            const int nblk = (T.nsegm + CHB_TILE_SB - 1) / CHB_TILE_SB, ngrp = std::max(1, 8 / nblk);
            const int TS = 32 * ngrp, SP = nblk * CHB_TILE_SB;
            const size_t arrt = (size_t)SP * TS * sizeof(cd);
            p->tile_threads = 32 * nblk * ngrp;
            p->tile_smem[0] = red_bytes + arrt;
            p->tile_smem[1] = red_bytes + 3 * arrt;
            p->tile_smem[2] = red_bytes + 4 * arrt;
            if (p->tile_smem[2] <= p->smem_limit) {
                S.pair_tilek = 1;
                S.tile_nblk = nblk;
                S.tile_ngrp = ngrp;
                S.pair_tile = TS;
                S.pair_ntiles = (segm_store + TS - 1) / TS;
            }
        }
    }
    p->nint = nint;
    p->ready = true;
    p->fused_active = 0;
    p->fused4_ok = 0;
    {
        int rcf = fused_setup(p, loops, segm_size);
        if (rcf < 0) { free_level(p); return chb_fail(CHB_ERR_CUDA, "chb_plan_set_nint: fused setup failed"); }
        if (T.vel) rcf = 0;  // velocity plans run the staged kernels (their scaling is not in the fused load passes)
        p->fused_active = rcf;
        p->fused5_active = 0;
        if (!rcf && !T.vel) {  // layouts versions 3 / 4 refuse: the general fused kernel
            int rc5 = fused5_setup(p, loops, segm_size, segm_store);
            if (rc5 < 0) { free_level(p); return chb_fail(CHB_ERR_CUDA, "chb_plan_set_nint: fused setup failed"); }
            if (getenv("CHB_FUSED5") && atoi(getenv("CHB_FUSED5")) == 0) rc5 = 0;  // A/B knob
            p->fused5_active = rc5;
            p->fused_active = rc5;
        }
        p->big_active = 0;
        {  // staged path (also the A/B path of fused layouts): parity-split transform kernels where the layout allows them
            static const int big_on = [] { const char *e = getenv("CHB_BIG"); return e ? atoi(e) : 1; }();    // A/B knobs, read once
            static const int big_r = [] { const char *e = getenv("CHB_BIG_R"); return e ? atoi(e) : 0; }();
            if (big_r) p->big_force_r = big_r;
            for (int i = 0; i < 3; ++i) {
                int rcb = big_on ? big_setup(p, loops, segm_size, segm_store, 2 << i, p->bigsys[i], &p->big_smem[i]) : 0;
                if (rcb < 0) { free_level(p); return chb_fail(CHB_ERR_CUDA, "chb_plan_set_nint: transform setup failed"); }
                p->big_ok[i] = rcb;
                p->big_min_len[i] = 1 << 30;
                for (int il = 0; rcb > 0 && il < T.nloop; ++il) p->big_min_len[i] = std::min(p->big_min_len[i], p->bigsys[i].loops[il].h.N);
                p->big_active |= rcb;
            }
        }
        p->ready = true;
    }
    if (p->vel_plan) {  // the sibling plan follows the level
        int rcv = chb_plan_set_nint(p->vel_plan, nint);
        if (rcv) return rcv;
    }
    return CHB_OK;
}

// velocity plan of params_to_segmvel (:6050-6118): the position plan with basis * i (:852-854), every loop on the GENERAL_SYM path
// (the velocity basis is never the identity, :1451-1466), segment and Adjust_after_last_gen rotations multiplied by their TimeRev
// (TimeDerivative, _ActionSym.pyx:268-291) and the scaling of changevar_mom_vel (:4830-4914)
static int ensure_vel_plan(chb_plan *p) {
    if (p->vel_plan) return CHB_OK;
    chb_plan *v = new chb_plan();
    v->device = p->device;
    v->T = p->T;
    HostTables &T = v->T;
    T.vel = 1;
    const int G = T.G;
    for (size_t j = 0; j + 1 < T.basis.size(); j += 2) {
        const double re = T.basis[j], im = T.basis[j + 1];
        T.basis[j] = -im;
        T.basis[j + 1] = re;
    }
    for (int il = 0; il < T.nloop; ++il) {
        T.shortcut[il] = CHB_GENERAL_SYM;
        for (int i = 0; i < G * G; ++i) T.alg_rot[(size_t)il * G * G + i] *= (double)T.alg_timerev[il];
    }
    for (int s = 0; s < T.nsegm; ++s) {
        for (int i = 0; i < G * G; ++i) T.seg_rot[(size_t)s * G * G + i] *= (double)T.seg_timerev[s];
        if (T.seg_timerev[s] < 0) T.seg_isid[s] = 0;
    }
    v->smem_limit = p->smem_limit;
    v->num_sms = p->num_sms;
    v->chunk_user = p->chunk_user;
    memset(&v->sys, 0, sizeof(v->sys));
    int rc = chb_plan_set_nint(v, p->nint);
    if (rc) { chb_plan_destroy(v); return rc; }
    p->vel_plan = v;
    return CHB_OK;
}

extern "C" int chb_plan_query(const chb_plan *p, int what, int64_t *v) {
    CHB_REQUIRE(p && v, CHB_ERR_INVALID_ARGUMENT, "chb_plan_query: NULL argument");
    if (what == CHB_Q_DEVICE) { *v = p->device; return CHB_OK; }
    if (what == CHB_Q_LAUNCHES) { *v = p->launches; return CHB_OK; }
    if (what == CHB_Q_NLOOP) { *v = p->T.nloop; return CHB_OK; }
    if (what == CHB_Q_NBIN) { *v = p->T.nbin; return CHB_OK; }
    if (what == CHB_Q_NSEGM) { *v = p->T.nsegm; return CHB_OK; }
    if (what == CHB_Q_GEODIM) { *v = p->T.G; return CHB_OK; }
    CHB_REQUIRE(p->ready, CHB_ERR_NOT_READY, "chb_plan_query: call chb_plan_set_nint first");
    switch (what) {
        case CHB_Q_NINT: *v = p->nint; break;
        case CHB_Q_NPARAMS: *v = p->sys.nparams; break;
        case CHB_Q_SEGM_SIZE: *v = p->sys.segm_size; break;
        case CHB_Q_SEGM_STORE: *v = p->sys.segm_store; break;
        case CHB_Q_SEGMPOS_SIZE: *v = p->sys.segsz; break;
        case CHB_Q_NPARAMS_INCL_O: *v = p->nparams_incl_o; break;
        case CHB_Q_STAGED_TRANSFORM: *v = p->big_active ? (p->big_ok[0] ? 2 : 0) + (p->big_ok[1] ? 4 : 0) + (p->big_ok[2] ? 8 : 0) : 0; break;
        case CHB_Q_PAIR_KERNEL: *v = p->sys.pair_tilek ? 2 : (p->sys.pair_dense ? 1 : 0); break;
#ifdef CHB_HOSTSIM
        case CHB_Q_HOST_POINTERS: *v = 1; break;
#else
        case CHB_Q_HOST_POINTERS: *v = 0; break;
#endif
        case CHB_Q_FUSED:
            *v = (p->fused_enabled && p->fused_active)
                     ? (p->fused5_active ? 5 : ((p->fused_enabled >= 4 && p->fused4_ok) ? 4 : std::min(p->fused_enabled, 3)))
                     : 0;
            break;
        case CHB_Q_CHUNK: {
            int64_t c = p->chunk_user;
            if (c <= 0) {
                long long per_state = std::max<long long>(1, p->sys.segsz * 8 * 3);
                c = std::max<long long>((48LL << 20) / per_state, (2LL * p->num_sms + p->T.nloop - 1) / p->T.nloop);
                c = std::max<int64_t>(c, 1);
            }
            *v = c;
            break;
        }
        default: return chb_fail(CHB_ERR_INVALID_ARGUMENT, "chb_plan_query: unknown query");
    }
    return CHB_OK;
}

extern "C" int chb_plan_set_chunk(chb_plan *p, int64_t chunk) {
    CHB_REQUIRE(p && chunk >= 0, CHB_ERR_INVALID_ARGUMENT, "chb_plan_set_chunk: bad argument");
    p->chunk_user = chunk;
    return CHB_OK;
}
extern "C" int chb_plan_set_fused(chb_plan *p, int enable) {
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, "chb_plan_set_fused: NULL plan");
    #ifdef CHF_KEEP_OLD
    p->fused_enabled = enable < 0 ? 0 : (enable > 4 ? 4 : enable);
#else
    p->fused_enabled = enable > 0 ? (enable >= 4 ? 4 : 3) : 0;  // the older fused variants (1, 2) are compiled only with -DCHF_KEEP_OLD
#endif
    return CHB_OK;
}

extern "C" int chb_plan_set_split(chb_plan *p, int split) {
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, "chb_plan_set_split: NULL plan");
    CHB_REQUIRE(split == 0 || split == 2 || split == 4 || split == 8, CHB_ERR_INVALID_ARGUMENT, "chb_plan_set_split: split must be 0, 2, 4 or 8");
    p->big_force_r = split;
    return CHB_OK;
}

// ------------------------------------------------------------------------------------------------
// workspace layout (per chunk of `nb` states)
// ------------------------------------------------------------------------------------------------
struct Workspace {
    double *segA, *segB, *segC;  // [nb][segsz]
    double *kin_part;            // [nb][nloop]
    double *pot_part;            // [nb][ntiles]
    cd *scratch;                 // [nb][scratch_per_state]
};
static size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }
static int64_t plan_chunk(const chb_plan *p, int64_t batch) {
    int64_t c = 1;
    chb_plan_query(p, CHB_Q_CHUNK, &c);
    return std::max<int64_t>(1, std::min<int64_t>(c, std::max<int64_t>(batch, 1)));
}
static size_t ws_bytes_for(const chb_plan *p, int64_t nb) {
    const ChbSys &S = p->sys;
    size_t b = 0;
    b += 3 * align_up((size_t)nb * S.segsz * sizeof(double));
    b += align_up((size_t)nb * S.nloop * sizeof(double));
    b += align_up((size_t)nb * S.pair_ntiles * sizeof(double));
    bool any_scratch = false;
    for (int il = 0; il < S.nloop; ++il)
        if (!p->loop_in_smem[il]) any_scratch = true;
    b += align_up(any_scratch ? (size_t)nb * S.scratch_per_state * sizeof(cd) : 16);
    return b;
}
static Workspace ws_carve(const chb_plan *p, void *base, int64_t nb) {
    const ChbSys &S = p->sys;
    Workspace w;
    char *q = (char *)base;
    size_t seg = align_up((size_t)nb * S.segsz * sizeof(double));
    w.segA = (double *)q; q += seg;
    w.segB = (double *)q; q += seg;
    w.segC = (double *)q; q += seg;
    w.kin_part = (double *)q; q += align_up((size_t)nb * S.nloop * sizeof(double));
    w.pot_part = (double *)q; q += align_up((size_t)nb * S.pair_ntiles * sizeof(double));
    w.scratch = (cd *)q;
    return w;
}

extern "C" int chb_workspace_bytes(const chb_plan *p, int64_t batch, size_t *out) {
    CHB_REQUIRE(p && out && batch >= 0, CHB_ERR_INVALID_ARGUMENT, "chb_workspace_bytes: bad argument");
    CHB_REQUIRE(p->ready, CHB_ERR_NOT_READY, "chb_workspace_bytes: call chb_plan_set_nint first");
    *out = ws_bytes_for(p, plan_chunk(p, batch)) + 256;
    if (p->vel_plan && p->vel_plan->ready)  // chb_params_to_segmvel_dev runs on the sibling plan with the same workspace
        *out = std::max(*out, ws_bytes_for(p->vel_plan, plan_chunk(p->vel_plan, batch)) + 256);
    return CHB_OK;
}

// ------------------------------------------------------------------------------------------------
// kernels (GPU) / block loops (hostsim)
// ------------------------------------------------------------------------------------------------
#ifndef CHB_HOSTSIM
// tensor-memory stash: ND doubles of this thread <-> columns [col, col + 2 ND) of its lane (warp-collective)
template <int ND>
__device__ __forceinline__ void dev_tmem_st(unsigned taddr, const double *v) {
#pragma unroll
    for (int i = 0; i < ND; i += 4) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr + 2 * i),
                     "r"(__double2loint(v[i])), "r"(__double2hiint(v[i])), "r"(__double2loint(v[i + 1])),
                     "r"(__double2hiint(v[i + 1])), "r"(__double2loint(v[i + 2])), "r"(__double2hiint(v[i + 2])),
                     "r"(__double2loint(v[i + 3])), "r"(__double2hiint(v[i + 3]))
                     : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
template <int ND>
__device__ __forceinline__ void dev_tmem_ld(unsigned taddr, double *v) {
    unsigned w[2 * ND];
#pragma unroll
    for (int i = 0; i < ND; i += 4) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=r"(w[2 * i]), "=r"(w[2 * i + 1]), "=r"(w[2 * i + 2]), "=r"(w[2 * i + 3]), "=r"(w[2 * i + 4]),
                       "=r"(w[2 * i + 5]), "=r"(w[2 * i + 6]), "=r"(w[2 * i + 7])
                     : "r"(taddr + 2 * i)
                     : "memory");
    }
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < ND; ++i) v[i] = __hiloint2double((int)w[2 * i + 1], (int)w[2 * i]);
}
struct DevGrp {
    int g, T;
    unsigned taddr;  // tensor-memory address of this thread's stash (lane quarter of its warp, column block of warp / 4)
    unsigned bar_base;  // shared-memory address of the CTA's mbarrier array (two per group: chunk A, chunk B)
    // ---- bulk asynchronous copies (cp.async.bulk) used by the version-4 fused kernel; called by ONE thread of the group
    // one mbarrier per (warp of the CTA, pass): the rows of the warp's butterflies of that pass
    __device__ __forceinline__ unsigned bar_addr(int warp, int it) const {
        return bar_base + 8u * (unsigned)(2 * (g * (T >> 5) + warp) + it);
    }
    __device__ __forceinline__ void tma_expect(int warp, int it, unsigned bytes) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_addr(warp, it)), "r"(bytes) : "memory");
    }
    __device__ __forceinline__ void tma_load(void *smem_dst, const void *gsrc, unsigned bytes, int warp, int it) {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         (unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc), "r"(bytes), "r"(bar_addr(warp, it)) : "memory");
    }
    // every thread: wait until the rows of its warp's butterflies of pass `it` have landed
    __device__ __forceinline__ void tma_wait(int it, unsigned parity) {
        unsigned done = 0;
        const unsigned a = bar_addr(((int)threadIdx.x - g * T) >> 5, it);
        while (!done) {
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(a), "r"(parity) : "memory");
        }
    }
    __device__ __forceinline__ void tma_store(void *gdst, const void *smem_src, unsigned bytes) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(
                         (unsigned)__cvta_generic_to_shared(smem_src)), "r"(bytes) : "memory");
    }
    __device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
    __device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
    __device__ __forceinline__ void tma_store_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
    __device__ __forceinline__ void fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
    __device__ __forceinline__ void warp_sync() { __syncwarp(); }
    // f(warp of the group, lane) on every thread, no barrier
    template <class F>
    __device__ __forceinline__ void lanes(F f) {
        const int tid = (int)threadIdx.x - g * T;
        f(tid >> 5, tid & 31);
    }
    // f1, warp barrier, f2: the threads of a warp exchange data through shared memory between the two halves
    template <class S, class F1, class F2>
    __device__ __forceinline__ void par2(F1 f1, F2 f2) {
        S s;
        const int tid = (int)threadIdx.x - g * T;
        f1(tid, T, s);
        __syncwarp();
        f2(tid, T, s);
    }
    template <int ND>
    __device__ __forceinline__ void tmem_st(int col, const double *v) { dev_tmem_st<ND>(taddr + (unsigned)col, v); }
    template <int ND>
    __device__ __forceinline__ void tmem_ld(int col, double *v) { dev_tmem_ld<ND>(taddr + (unsigned)col, v); }
    template <class S, int NX, class F1, class F2, class F3>
    __device__ __forceinline__ void par_xchg2(F1 f1, F2 f2, F3 f3) {
        S s;
        const int tid = (int)threadIdx.x - g * T;
        f1(tid, T, s);
#pragma unroll
        for (int i = 0; i < NX; ++i) {
            s.x[i].x = __shfl_xor_sync(0xffffffffu, s.x[i].x, s.xm);
            s.x[i].y = __shfl_xor_sync(0xffffffffu, s.x[i].y, s.xm);
        }
        f2(tid, T, s);
#pragma unroll
        for (int i = 0; i < NX; ++i) {
            s.x[i].x = __shfl_xor_sync(0xffffffffu, s.x[i].x, s.xm);
            s.x[i].y = __shfl_xor_sync(0xffffffffu, s.x[i].y, s.xm);
        }
        f3(tid, T, s);
    }
    // phase of one thread group: the group's own named barrier, so groups drift apart and overlap each other
    template <class F>
    __device__ __forceinline__ void par(F f) {
        f((int)threadIdx.x - g * T, T);
        asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(T) : "memory");
    }
    __device__ __forceinline__ void sync() { asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(T) : "memory"); }
    // f1, exchange of s.x[0..NX) with lane ^ s.xm (0 keeps the own values), f2; no barrier.  T is a multiple of 32
    // and every thread of the group makes the call, so all lanes of a warp take part in the shuffles.
    template <class S, int NX, class F1, class F2>
    __device__ __forceinline__ void par_xchg(F1 f1, F2 f2) {
        S s;
        const int tid = (int)threadIdx.x - g * T;
        f1(tid, T, s);
#pragma unroll
        for (int i = 0; i < NX; ++i) {
            s.x[i].x = __shfl_xor_sync(0xffffffffu, s.x[i].x, s.xm);
            s.x[i].y = __shfl_xor_sync(0xffffffffu, s.x[i].y, s.xm);
        }
        f2(tid, T, s);
    }
    template <class F>
    __device__ __forceinline__ void one(F f) {
        if ((int)threadIdx.x == g * T) f();
    }
};
struct DevCtx {
    unsigned tmem_thread;
    unsigned bar_base;
    // sum of red[0 .. blockDim) by the first warp (strided partial sums, then a shuffle tree), stored by thread 0; NO barrier
    __device__ __forceinline__ void reduce_store(const double *red, double *out) {
        if (threadIdx.x < 32) {
            double s = 0.0;
            for (int j = (int)threadIdx.x; j < (int)blockDim.x; j += 32) s += red[j];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (threadIdx.x == 0) *out = s;
        }
    }
    template <class F>
    __device__ __forceinline__ void par(F f) {
        f((int)threadIdx.x, (int)blockDim.x);
        __syncthreads();
    }
    // blockDim.x == ngroups * T
    template <class F>
    __device__ __forceinline__ void groups(int ngroups, int T, F f) {
        (void)ngroups;
        DevGrp grp{(int)threadIdx.x / T, T, tmem_thread, bar_base};
        f(grp);
        __syncthreads();
    }
};
extern __shared__ __align__(16) unsigned char chb_smem_raw[];
#define CHB_KERNEL __global__
#define CHB_BLOCK_ID ((long long)blockIdx.x)
#define CHB_NTHREADS ((int)blockDim.x)
#define CHB_CTX_DECL DevCtx ctx{0u, 0u}
#define CHB_GRID_DIM ((long long)gridDim.x)
#define CHB_SMEM chb_smem_raw
#else
struct HostGrp {
    int g, T;
    std::vector<double> *tmem;  // emulated tensor-memory stash, CHF_MAX_THREADS x 256 doubles
    int tmem_tid;               // thread currently running (set by par_xchg2)
    template <int ND>
    void tmem_st(int col, const double *v) {
        for (int i = 0; i < ND; ++i) (*tmem)[(size_t)(g * T + tmem_tid) * 256 + col / 2 + i] = v[i];
    }
    template <int ND>
    void tmem_ld(int col, double *v) {
        for (int i = 0; i < ND; ++i) v[i] = (*tmem)[(size_t)(g * T + tmem_tid) * 256 + col / 2 + i];
    }
    template <class S, int NX, class F1, class F2, class F3>
    void par_xchg2(F1 f1, F2 f2, F3 f3) {
        std::vector<S> st((size_t)T);
        auto xchg = [&]() {
            for (int t = 0; t < T; ++t) {
                const int u = t ^ st[t].xm;
                if (u > t)
                    for (int i = 0; i < NX; ++i) std::swap(st[t].x[i], st[u].x[i]);
            }
        };
        for (int t = 0; t < T; ++t) { tmem_tid = t; f1(t, T, st[t]); }
        xchg();
        for (int t = 0; t < T; ++t) { tmem_tid = t; f2(t, T, st[t]); }
        xchg();
        for (int t = 0; t < T; ++t) { tmem_tid = t; f3(t, T, st[t]); }
    }
    template <class F>
    void par(F f) {
        for (int t = 0; t < T; ++t) f(t, T);
    }
    void sync() {}
    template <class S, int NX, class F1, class F2>
    void par_xchg(F1 f1, F2 f2) {
        std::vector<S> st((size_t)T);
        for (int t = 0; t < T; ++t) { tmem_tid = t; f1(t, T, st[t]); }
        for (int t = 0; t < T; ++t) {
            const int u = t ^ st[t].xm;
            if (u > t)
                for (int i = 0; i < NX; ++i) std::swap(st[t].x[i], st[u].x[i]);
        }
        for (int t = 0; t < T; ++t) { tmem_tid = t; f2(t, T, st[t]); }
    }
    template <class F>
    void one(F f) { f(); }
    // bulk copies complete at once in the emulation
    void tma_expect(int, int, unsigned) {}
    void tma_load(void *dst, const void *src, unsigned bytes, int, int) { memcpy(dst, src, bytes); }
    void tma_wait(int, unsigned) {}
    void warp_sync() {}
    template <class F>
    void lanes(F f) {
        for (int t = 0; t < T; ++t) f(t >> 5, t & 31);
    }
    void tma_store(void *dst, const void *src, unsigned bytes) { memcpy(dst, src, bytes); }
    void tma_store_commit() {}
    void tma_store_wait_read() {}
    void tma_store_wait_read1() {}
    void fence_async() {}
    template <class S, class F1, class F2>
    void par2(F1 f1, F2 f2) {
        std::vector<S> st((size_t)T);
        for (int t = 0; t < T; ++t) f1(t, T, st[t]);
        for (int t = 0; t < T; ++t) f2(t, T, st[t]);
    }
};
struct HostCtx {
    int nthr;
    std::vector<double> tmem;  // emulated tensor memory, sized by k_fused
    void reduce_store(const double *red, double *out) {  // same association as the device version
        double part[32];
        for (int l = 0; l < 32; ++l) {
            part[l] = 0.0;
            for (int j = l; j < nthr; j += 32) part[l] += red[j];
        }
        for (int o = 16; o > 0; o >>= 1)
            for (int l = 0; l < o; ++l) part[l] += part[l + o];
        *out = part[0];
    }
    template <class F>
    void par(F f) {
        for (int t = 0; t < nthr; ++t) f(t, nthr);
    }
    template <class F>
    void groups(int ngroups, int T, F f) {
        for (int g = 0; g < ngroups; ++g) {
            HostGrp grp{g, T, &tmem, 0};
            f(grp);
        }
    }
};
static thread_local long long sim_block_id;
static thread_local long long sim_grid_dim;
static thread_local int sim_nthreads;
static thread_local unsigned char *sim_smem;
#define CHB_KERNEL static
#define CHB_BLOCK_ID sim_block_id
#define CHB_GRID_DIM sim_grid_dim
#define CHB_NTHREADS sim_nthreads
#define CHB_CTX_DECL HostCtx ctx{sim_nthreads}
#define CHB_SMEM sim_smem
#endif

// one CTA per (state, loop): params -> segmpos (+ kinetic partial)
CHB_KERNEL void k_fwd(ChbSys S, const double *params, double *segmpos, cd *scratch, double *kin_part,
                      unsigned long long smem_mask) {
    CHB_CTX_DECL;
    const long long blk = CHB_BLOCK_ID;
    const int il = (int)(blk % S.nloop);
    const long long b = blk / S.nloop;
    if (S.skip && S.skip[b]) return;
    const ChbLoop &L = S.loops[il];
    double *red = (double *)CHB_SMEM;
    cd *buf = ((smem_mask >> il) & 1ULL) ? (cd *)(CHB_SMEM + sizeof(double) * 1024)
                                         : scratch + b * S.scratch_per_state + L.scratch_off;
    cta_params_to_segmpos(ctx, S, L, params + b * S.nparams, segmpos ? segmpos + b * S.segsz : (double *)0, buf, red,
                          kin_part ? kin_part + b * S.nloop + il : (double *)0);
}

// one CTA per (state, loop): segment array -> params (mode 0 transpose, 1 inverse)
CHB_KERNEL void k_adj(ChbSys S, const double *seg, double *out, const double *kin_src, cd *scratch, int mode,
                      unsigned long long smem_mask) {
    CHB_CTX_DECL;
    const long long blk = CHB_BLOCK_ID;
    const int il = (int)(blk % S.nloop);
    const long long b = blk / S.nloop;
    if (S.skip && S.skip[b]) return;
    const ChbLoop &L = S.loops[il];
    cd *buf = ((smem_mask >> il) & 1ULL) ? (cd *)(CHB_SMEM + sizeof(double) * 1024)
                                         : scratch + b * S.scratch_per_state + L.scratch_off;
    cta_segm_to_params(ctx, S, L, seg + b * S.segsz, out + b * S.nparams, kin_src ? kin_src + b * S.nparams : (const double *)0,
                       buf, mode);
}

template <int MODE>
CHB_KERNEL void k_pair(ChbSys S, const double *seg, const double *dseg, double *out, double *pot_part, double globalmul) {
    CHB_CTX_DECL;
    const long long blk = CHB_BLOCK_ID;
    const int tile = (int)(blk % S.pair_ntiles);
    const long long b = blk / S.pair_ntiles;
    if (S.skip && S.skip[b]) return;
    double *red = (double *)CHB_SMEM;
    const size_t arr = (size_t)S.nsegm * S.G * S.pair_tile;
    double *spos = red + 1024;
    double *sout = spos + arr;
    double *sdpos = sout + arr;
    cta_pair<MODE>(ctx, S, seg + b * S.segsz, dseg ? dseg + b * S.segsz : (const double *)0,
                   out ? out + b * S.segsz : (double *)0, pot_part ? pot_part + b * S.pair_ntiles + tile : (double *)0, tile,
                   globalmul, spos, sdpos, sout, red);
}

template <int MODE, bool UNI, bool LAW0>
CHB_KERNEL void k_pair_dense(ChbSys S, const double *seg, const double *dseg, double *out, double *pot_part, double globalmul) {
    CHB_CTX_DECL;
    const long long blk = CHB_BLOCK_ID;
    const int tile = (int)(blk % S.pair_ntiles);
    const long long b = blk / S.pair_ntiles;
    if (S.skip && S.skip[b]) return;
    double *red = (double *)CHB_SMEM;
    double *sw = red + 1024;
    cd *spos = (cd *)(sw + (UNI ? (size_t)0 : (size_t)S.nsegm * S.pair_spad));
    cd *sout = spos + (size_t)S.pair_tile * S.pair_spad;
    cd *sdpos = sout + (size_t)S.pair_tile * S.pair_spad;
    cta_pair_dense<MODE, UNI, LAW0>(ctx, S, seg + b * S.segsz, dseg ? dseg + b * S.segsz : (const double *)0,
                         out ? out + b * S.segsz : (double *)0, pot_part ? pot_part + b * S.pair_ntiles + tile : (double *)0, tile,
                         globalmul, CHB_NTHREADS, spos, sdpos, sout, sw, red);
}

template <int MODE, bool UNI, bool LAW0>
CHB_KERNEL void
#ifndef CHB_HOSTSIM
__launch_bounds__(256, 2)
#endif
k_pair_tile(ChbSys S, const double *seg, const double *dseg, double *out, double *pot_part, double globalmul) {
    CHB_CTX_DECL;
    const long long blk = CHB_BLOCK_ID;
    const int tile = (int)(blk % S.pair_ntiles);
    const long long b = blk / S.pair_ntiles;
    if (S.skip && S.skip[b]) return;
    double *red = (double *)CHB_SMEM;
    const size_t arr = (size_t)S.tile_nblk * CHB_TILE_SB * 32 * S.tile_ngrp;
    cd *xs = (cd *)(red + 1024);
    cd *accA = xs + arr, *accB = accA + arr, *dxs = accB + arr;
    cta_pair_tile<MODE, UNI, LAW0>(ctx, S, seg + b * S.segsz, dseg ? dseg + b * S.segsz : (const double *)0,
                        out ? out + b * S.segsz : (double *)0, pot_part ? pot_part + b * S.pair_ntiles + tile : (double *)0, tile,
                        globalmul, S.tile_nblk, S.tile_ngrp, xs, dxs, accA, accB, red);
}

// A/B variants of the tiled kernel for energy + gradient, uniform weights, closed-form law (knob CHB_TILE_VAR, read once):
// 1: partner sub-blocks of 2 segments (the default since the A/B), 2: sub-blocks of 8, 3: sub-blocks of 4 -- 2 and 3 with one CTA per SM
// and up to 255 registers
#ifndef CHB_HOSTSIM
template <int SBBV, int MINB>
__global__ void __launch_bounds__(256, MINB)
k_pair_tile_var(ChbSys S, const double *seg, const double *dseg, double *out, double *pot_part, double globalmul) {
    CHB_CTX_DECL;
    const long long blk = CHB_BLOCK_ID;
    const int tile = (int)(blk % S.pair_ntiles);
    const long long b = blk / S.pair_ntiles;
    if (S.skip && S.skip[b]) return;
    double *red = (double *)CHB_SMEM;
    const size_t arr = (size_t)S.tile_nblk * CHB_TILE_SB * 32 * S.tile_ngrp;
    cd *xs = (cd *)(red + 1024);
    cd *accA = xs + arr, *accB = accA + arr, *dxs = accB + arr;
    cta_pair_tile<1, true, true, SBBV>(ctx, S, seg + b * S.segsz, dseg ? dseg + b * S.segsz : (const double *)0,
                                       out ? out + b * S.segsz : (double *)0, pot_part ? pot_part + b * S.pair_ntiles + tile : (double *)0,
                                       tile, globalmul, S.tile_nblk, S.tile_ngrp, xs, dxs, accA, accB, red);
}
#endif

CHB_KERNEL void k_kin(ChbSys S, const double *params, double *kin_out, int kin_stride) {
    CHB_CTX_DECL;
    const long long b = CHB_BLOCK_ID;
    cta_kin(ctx, S, params + b * S.nparams, kin_out + b * kin_stride, (double *)CHB_SMEM);
}

// out[b] = sum kin_part[b][0..nk) + psign * sum pot_part[b][0..np)
CHB_KERNEL void k_finish(double *out, const double *kin_part, int nk, const double *pot_part, int np, long long nstates,
                         double psign) {
#ifndef CHB_HOSTSIM
    long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b < nstates) out[b] = chb_finish_action(kin_part + b * nk, kin_part ? nk : 0, pot_part + b * np, pot_part ? np : 0, psign);
#else
    for (int t = 0; t < sim_nthreads; ++t) {
        long long b = sim_block_id * sim_nthreads + t;
        if (b < nstates) out[b] = chb_finish_action(kin_part + b * nk, kin_part ? nk : 0, pot_part + b * np, pot_part ? np : 0, psign);
    }
#endif
}


// one CTA per state: the whole evaluation in shared memory (chb_fused.cuh)
#ifndef CHB_HOSTSIM
#define CHF_BOUNDS __launch_bounds__(CHF_MAX_THREADS, 1)
#else
#define CHF_BOUNDS
#endif
template <int NSEG, int PMODE>
CHB_KERNEL void CHF_BOUNDS k_fused(ChfSys F, ChfArgs A) {
    CHB_CTX_DECL;
    const bool use_tmem = F.ver >= 3 && F.tmem_cols > 0;
#ifndef CHB_HOSTSIM
    __shared__ unsigned tmem_base_s;
    unsigned tmem_base = 0;
    if (use_tmem) {  // one warp allocates the columns, everybody reads the base address after a fenced barrier
        if (threadIdx.x < 32) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                             (unsigned)__cvta_generic_to_shared(&tmem_base_s)), "r"((unsigned)F.tmem_cols) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        tmem_base = tmem_base_s;
        const unsigned warp = threadIdx.x >> 5;
        ctx.tmem_thread = tmem_base + (((warp & 3u) * 32u) << 16) + (warp >> 2) * (unsigned)F.tmem_thread_cols;
    }
#else
    if (use_tmem) ctx.tmem.assign((size_t)CHF_MAX_THREADS * 256, 0.0);
#endif
    cta_fused_prologue(ctx, F, (cd *)(CHB_SMEM + CHF_RED_BYTES));
    // persistent CTAs: the grid is one resident wave, every CTA walks the batch with the grid stride
    // or, with A.live_list, over the states it names (written by the previous GMRES step); b_pf: the state whose rows are
    // prefetched into L2 (the one this CTA or a neighbour processes one wave later)
    const long long nwork = A.live_list ? (long long)*A.live_count : A.nstates;
    for (long long k = CHB_BLOCK_ID; k < nwork; k += CHB_GRID_DIM) {
        const long long b = A.live_list ? (long long)A.live_list[k] : k;
        if (A.skip && A.skip[b]) continue;  // CTA-uniform
        const long long kp = k + F.pf_stride;
        const long long b_pf = kp < nwork ? (A.live_list ? (long long)A.live_list[kp] : kp) : A.nstates;
        cta_fused<NSEG, PMODE>(ctx, F, A, b, b_pf, (double *)CHB_SMEM, (cd *)(CHB_SMEM + CHF_RED_BYTES));
    }
#ifndef CHB_HOSTSIM
    if (use_tmem) {
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (threadIdx.x < 32)
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((unsigned)F.tmem_cols) : "memory");
    }
#endif
}

// version 4 of the fused kernel (chb_fused4.cuh): persistent CTAs, rows moved by bulk asynchronous copies
// SEL: 0 = every state of the batch, grid stride; 1 = the states selected by A.skip / A.live_list (GMRES matvec)
template <int NSEG, int PMODE, int SEL = 0>
CHB_KERNEL void CHF_BOUNDS k_fused4(ChfSys F, ChfArgs A) {
    CHB_CTX_DECL;
#ifndef CHB_HOSTSIM
    __shared__ unsigned tmem_base_s;
    __shared__ __align__(8) unsigned long long load_bars[2 * (CHF_MAX_THREADS / 32)];
    unsigned tmem_base = 0;
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         (unsigned)__cvta_generic_to_shared(&tmem_base_s)), "r"((unsigned)F.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2 * (CHF_MAX_THREADS / 32); ++i)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((unsigned)__cvta_generic_to_shared(&load_bars[i])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    tmem_base = tmem_base_s;
    {
        const unsigned warp = threadIdx.x >> 5;
        ctx.tmem_thread = tmem_base + (((warp & 3u) * 32u) << 16) + (warp >> 2) * (unsigned)F.tmem_thread_cols;
        ctx.bar_base = (unsigned)__cvta_generic_to_shared(&load_bars[0]);
    }
    if (F.stagger_ns > 0 && A.nstates > (long long)gridDim.x) {
        // all CTAs do identical work: without a stagger their copy phases coincide and the memory system sees bursts.
        // CTA i starts i / gridDim of one state period late (the CTAs with the higher indices own fewer states).
        if (threadIdx.x == 0) {
            unsigned long long t0, t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            const unsigned long long wait = (unsigned long long)F.stagger_ns * blockIdx.x / gridDim.x;
            do {
                __nanosleep(500);
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            } while (t1 - t0 < wait);
        }
        __syncthreads();
    }
#else
    ctx.tmem.assign((size_t)CHF_MAX_THREADS * 256, 0.0);
#endif
    double *red = (double *)CHB_SMEM;
    cd *sm = (cd *)(CHB_SMEM + CHF_RED_BYTES);
    // the states of this CTA: grid stride over the batch (the ones flagged in A.skip left out) or over A.live_list; the next
    // state is known one state ahead: its rows are requested while the current one drains
    unsigned parity = 0;
    double *red_alt = (double *)(sm + F.tw4_off + 256);  // second buffer of partial sums, behind the twiddle tables
    if (SEL == 0) {
        cta_fused4_first_loads(ctx, F, A, CHB_BLOCK_ID, sm);
        cta_fused4_prologue(ctx, F, sm);
        for (long long b = CHB_BLOCK_ID; b < A.nstates; b += CHB_GRID_DIM, parity ^= 1u)
            cta_fused4<NSEG, PMODE>(ctx, F, A, b, b + CHB_GRID_DIM, parity, parity ? red_alt : red, sm);
    } else {
        // (32-bit positions, the list length re-read from memory: the loop state stays out of the register budget of the passes)
        auto settle = [&](int k) {  // first position >= k (same CTA) that is not flagged
            if (!A.live_list)
                while (A.skip && k < (int)A.nstates && A.skip[k]) k += (int)CHB_GRID_DIM;
            return k;
        };
        auto state_at = [&](int k) -> long long {  // position -> state (nstates: none left)
            if (A.live_list) return k < *A.live_count ? (long long)A.live_list[k] : A.nstates;
            return k < (int)A.nstates ? (long long)k : A.nstates;
        };
        int k = settle((int)CHB_BLOCK_ID);
        const long long b_first = state_at(k);
        cta_fused4_first_loads(ctx, F, A, b_first, sm);
        cta_fused4_prologue(ctx, F, sm);
        for (long long b = b_first; b < A.nstates; parity ^= 1u) {
            k = settle(k + (int)CHB_GRID_DIM);
            const long long bn = state_at(k);
            cta_fused4<NSEG, PMODE>(ctx, F, A, b, bn, parity, parity ? red_alt : red, sm);
            b = bn;
        }
    }
#ifndef CHB_HOSTSIM
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // the result rows of the last state have left shared memory
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((unsigned)F.tmem_cols) : "memory");
#endif
}


// general fused kernel (chb_fused5.cuh): persistent CTAs, one state at a time per CTA, any mix of RFFT / GENERAL loops
#ifndef CHB_HOSTSIM
#define CH5_SYS_PARAM const __grid_constant__ Ch5Sys F
#else
#define CH5_SYS_PARAM const Ch5Sys &F
#endif
template <int NSEG, int PMODE>
CHB_KERNEL void CHF_BOUNDS k_fused5(CH5_SYS_PARAM, ChfArgs A) {
    CHB_CTX_DECL;
    double *red = (double *)CHB_SMEM;
    cd *bnd = (cd *)(CHB_SMEM + sizeof(double) * CHF_MAX_THREADS);  // 1 KB behind the partial sums
    cd *sm = (cd *)(CHB_SMEM + CHF_RED_BYTES);
    cta_fused5_prologue(ctx, F, sm);
    const long long nwork = A.live_list ? (long long)*A.live_count : A.nstates;
    for (long long k = CHB_BLOCK_ID; k < nwork; k += CHB_GRID_DIM) {
        const long long b = A.live_list ? (long long)A.live_list[k] : k;
        if (A.skip && A.skip[b]) continue;  // CTA-uniform
        const long long kp = k + CHB_GRID_DIM;
        const long long b_pf = kp < nwork ? (A.live_list ? (long long)A.live_list[kp] : kp) : A.nstates;
        cta_fused5<NSEG, PMODE>(ctx, F, A, b, b_pf, red, sm, bnd);
    }
}

// residue-split staged transforms (chb_big.cuh): one CTA per (state, loop, residue modulo R)
#ifndef CHB_HOSTSIM
#define CHB_BIG_BOUNDS __launch_bounds__(CHB_BIG_MAX_THREADS, 1)
#else
#define CHB_BIG_BOUNDS
#endif
// With one CTA per SM all CTAs of a wave run their phases in lock-step: every SM reads parameter rows at the same time (an HBM
// burst), then every SM computes.  A start offset for the first-wave CTAs on odd SMs shifts half of the chip by about one phase, so
// that the memory phase of one half overlaps the FFT phase of the other; later CTAs inherit the shift (knob CHB_BIG_STAGGER_NS).
CHB_D void chb_big_stagger(const ChbBigSys &F, long long blk) {
#ifndef CHB_HOSTSIM
    if (F.stagger_ns > 0 && blk < F.first_wave) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        if (smid & 1u) {
            for (int left = F.stagger_ns; left > 0; left -= 4000) __nanosleep(left < 4000 ? left : 4000);
        }
    }
#else
    (void)F; (void)blk;
#endif
}
template <int R>
CHB_KERNEL void CHB_BIG_BOUNDS k_big_fwd(ChbBigSys F, ChbSys S, const double *params, double *segmpos, double *kin_part) {
    CHB_CTX_DECL;
    const long long blk = CHB_BLOCK_ID;
    const int q = (int)(blk % R);
    const int il = (int)((blk / R) % F.nloop);
    const long long b = (blk / R) / F.nloop;
    if (S.skip && S.skip[b]) return;
    const ChbBigLoop &L = F.loops[il];
    double *red = (double *)CHB_SMEM;
    cd *buf = (cd *)(CHB_SMEM + CHF_RED_BYTES);
    cd *tw = buf + L.h.N;
    ChbBigSlot *slots = (ChbBigSlot *)(tw + F.tw_smem_len);
    chb_big_stagger(F, blk);
    cta_big_prologue(ctx, F, S, L, tw, slots);
    cta_big_fwd<R>(ctx, F, L, q, params + b * F.nparams, segmpos ? segmpos + b * F.segsz : (double *)0,
                   kin_part ? kin_part + b * F.nloop + il : (double *)0, buf, tw, slots, red);
}
template <int R>
CHB_KERNEL void CHB_BIG_BOUNDS k_big_adj(ChbBigSys F, ChbSys S, const double *seg, double *out, const double *kin_src) {
    CHB_CTX_DECL;
    const long long blk = CHB_BLOCK_ID;
    const int q = (int)(blk % R);
    const int il = (int)((blk / R) % F.nloop);
    const long long b = (blk / R) / F.nloop;
    if (S.skip && S.skip[b]) return;
    const ChbBigLoop &L = F.loops[il];
    cd *buf = (cd *)(CHB_SMEM + CHF_RED_BYTES);
    cd *tw = buf + L.h.N;
    ChbBigSlot *slots = (ChbBigSlot *)(tw + F.tw_smem_len);
    chb_big_stagger(F, blk);
    cta_big_prologue(ctx, F, S, L, tw, slots);
    cta_big_adj<R>(ctx, F, L, q, seg + b * F.segsz, out + b * F.nparams, kin_src ? kin_src + b * F.nparams : (const double *)0, buf, tw, slots);
}

#ifndef CHB_HOSTSIM
// Opt-in to large dynamic shared memory: once per launch site and device, to the device maximum minus the kernel's static shared
// memory (a constant, so host threads that launch the same kernel with different sizes do not race on the attribute).
template <class K>
static cudaError_t chb_smem_opt_in(K kern, size_t limit) {
    cudaFuncAttributes fa;
    cudaError_t e = cudaFuncGetAttributes(&fa, kern);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)limit - (int)fa.sharedSizeBytes);
}
#define CHB_LAUNCH(p, kern, grid, block, smem, stream, ...)                                                  \
    do {                                                                                                     \
        if ((smem) > 32 * 1024) { /* static shared memory counts towards the 48 KB default limit too */    \
            static std::atomic<unsigned long long> opted_{0};                                                \
            const unsigned long long bit_ = 1ULL << ((p)->device & 63);                                      \
            if (!(opted_.load(std::memory_order_acquire) & bit_)) {                                          \
                CHB_CUDA(chb_smem_opt_in(kern, (p)->smem_limit));                                            \
                opted_.fetch_or(bit_, std::memory_order_release);                                            \
            }                                                                                                \
        }                                                                                                    \
        kern<<<(unsigned)(grid), (unsigned)(block), (smem), (stream)>>>(__VA_ARGS__);                       \
        CHB_CUDA(cudaGetLastError());                                                                        \
        (p)->launches++;                                                                                     \
    } while (0)
#else
#define CHB_LAUNCH(p, kern, grid, block, smem, stream, ...)             \
    do {                                                                \
        std::vector<unsigned char> smem_buf_((smem) + 64);              \
        sim_smem = smem_buf_.data();                                    \
        sim_smem += (16 - ((uintptr_t)sim_smem & 15)) & 15;             \
        sim_nthreads = (int)(block);                                    \
        sim_grid_dim = (long long)(grid);                               \
        for (long long blk_ = 0; blk_ < (long long)(grid); ++blk_) {    \
            sim_block_id = blk_;                                        \
            kern(__VA_ARGS__);                                          \
        }                                                               \
        (p)->launches++;                                                \
    } while (0)
#endif

static unsigned long long smem_mask(const chb_plan *p) {
    unsigned long long m = 0;
    for (int il = 0; il < p->sys.nloop && il < 64; ++il)
        if (p->loop_in_smem[il]) m |= 1ULL << il;
    return m;
}

// ------------------------------------------------------------------------------------------------
// operator drivers
// ------------------------------------------------------------------------------------------------
#define CHB_OP_PROLOGUE(name)                                                                                   \
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, name ": NULL plan");                                               \
    CHB_REQUIRE(p->ready, CHB_ERR_NOT_READY, name ": call chb_plan_set_nint first");                            \
    CHB_REQUIRE(batch >= 0, CHB_ERR_INVALID_ARGUMENT, name ": negative batch");                                 \
    if (batch == 0) return CHB_OK;                                                                              \
    CHB_REQUIRE(p->sys.nloop <= 64, CHB_ERR_UNSUPPORTED, name ": more than 64 loops");                          \
    const ChbSys &S = p->sys;                                                                                   \
    const int64_t chunk = plan_chunk(p, batch);                                                                 \
    CHB_REQUIRE(ws && ws_bytes >= ws_bytes_for(p, chunk), CHB_ERR_WORKSPACE, name ": workspace missing or too small"); \
    chb_stream_t st = (chb_stream_t)stream;                                                                     \
    void *ws_al = (void *)(((uintptr_t)ws + 255) & ~(uintptr_t)255);                                            \
    Workspace W = ws_carve(p, ws_al, chunk);                                                                    \
    const unsigned long long mask = smem_mask(p);                                                               \
    (void)W; (void)mask; (void)st; (void)S;
#ifndef CHB_HOSTSIM
#define CHB_SET_DEVICE() CHB_CUDA(cudaSetDevice(p->device))
#else
#define CHB_SET_DEVICE() \
    do {                 \
    } while (0)
#endif

// the system descriptor of a staged launch: inside the Newton driver it carries the byte mask of the states to leave out
static ChbSys stage_sys(const chb_plan *p, const ChbSys &S) {
    ChbSys Sy = S;
    Sy.skip = p->stage_skip ? p->stage_skip + p->stage_c0 : nullptr;
    return Sy;
}
// which residue split serves a launch of nb states: the smallest that exists when the batch fills the GPU (least folding work),
// a larger one while the CTAs would not (few live states at the end of a search: parallelism inside the state)
static int big_pick(const chb_plan *p, int64_t nb) {
    if (p->big_force_r)
        for (int i = 0; i < 3; ++i)
            if ((2 << i) == p->big_force_r && p->big_ok[i]) return i;
    int best = -1;
    for (int i = 0; i < 3 && best < 0; ++i)
        if (p->big_ok[i]) best = i;
    const long long want = 2LL * p->num_sms;
    for (int i = best + 1; i < 3; ++i) {
        if (!p->big_ok[i]) continue;
        if ((long long)nb * p->sys.nloop * (2 << best) >= want) break;
        if (p->big_min_len[i] < 512) break;  // shorter transforms do not pay for the folding pass
        best = i;
    }
    return best;
}
static int launch_fwd(chb_plan *p, int64_t nb, const double *params, double *segmpos, Workspace &W, bool want_kin,
                      unsigned long long mask, chb_stream_t st) {
    const ChbSys Sy = stage_sys(p, p->sys);
    if (p->big_active) {
        const int i = big_pick(p, nb);
        const ChbBigSys &F = p->bigsys[i];
        double *kp = want_kin ? W.kin_part : (double *)0;
        if (i == 0) CHB_LAUNCH(p, k_big_fwd<2>, 2 * nb * p->sys.nloop, F.T, p->big_smem[i], st, F, Sy, params, segmpos, kp);
        if (i == 1) CHB_LAUNCH(p, k_big_fwd<4>, 4 * nb * p->sys.nloop, F.T, p->big_smem[i], st, F, Sy, params, segmpos, kp);
        if (i == 2) CHB_LAUNCH(p, k_big_fwd<8>, 8 * nb * p->sys.nloop, F.T, p->big_smem[i], st, F, Sy, params, segmpos, kp);
        return CHB_OK;
    }
    CHB_LAUNCH(p, k_fwd, nb * p->sys.nloop, p->fft_threads, p->fft_smem, st, Sy, params, segmpos, W.scratch,
               want_kin ? W.kin_part : (double *)0, mask);
    return CHB_OK;
}
static int launch_adj(chb_plan *p, int64_t nb, const double *seg, double *out, const double *kin_src, Workspace &W,
                      int mode, unsigned long long mask, chb_stream_t st) {
    const ChbSys Sy = stage_sys(p, p->sys);
    if (p->big_active && mode == 0) {
        const int i = big_pick(p, nb);
        const ChbBigSys &F = p->bigsys[i];
        if (i == 0) CHB_LAUNCH(p, k_big_adj<2>, 2 * nb * p->sys.nloop, F.T, p->big_smem[i], st, F, Sy, seg, out, kin_src);
        if (i == 1) CHB_LAUNCH(p, k_big_adj<4>, 4 * nb * p->sys.nloop, F.T, p->big_smem[i], st, F, Sy, seg, out, kin_src);
        if (i == 2) CHB_LAUNCH(p, k_big_adj<8>, 8 * nb * p->sys.nloop, F.T, p->big_smem[i], st, F, Sy, seg, out, kin_src);
        return CHB_OK;
    }
    CHB_LAUNCH(p, k_adj, nb * p->sys.nloop, p->fft_threads, p->fft_smem, st, Sy, seg, out, kin_src, W.scratch, mode,
               mask);
    return CHB_OK;
}
static int launch_pair_sys(chb_plan *p, const ChbSys &Sy, int mode, int64_t nb, const double *seg, const double *dseg,
                           double *out, double *pot_part, double globalmul, chb_stream_t st);
static int launch_pair(chb_plan *p, int mode, int64_t nb, const double *seg, const double *dseg, double *out,
                       double *pot_part, double globalmul, chb_stream_t st) {
    return launch_pair_sys(p, p->sys, mode, nb, seg, dseg, out, pot_part, globalmul, st);
}
// Sy: the plan's system, possibly with another interaction law (the hash sums of chb_segmpos_to_hash_dev)
static int launch_pair_sys(chb_plan *p, const ChbSys &Sy_in, int mode, int64_t nb, const double *seg, const double *dseg,
                           double *out, double *pot_part, double globalmul, chb_stream_t st) {
    const ChbSys Sy = stage_sys(p, Sy_in);
    const long long grid = nb * p->sys.pair_ntiles;
    // uniform-weight variants only for the plan's own law and weights (the hash sums reuse the same weights)
    const bool uni = p->sys.pair_uniform != 0;
    const bool law0 = Sy.law_kind == 0;  // of THIS launch (the hash sums run other laws through the same kernels)
    if (p->sys.pair_tilek) {
        const bool tu = uni && (p->sys.nsegm % CHB_TILE_SB == 0);
#define CHB_TILE_LAUNCH2(M, U, W) CHB_LAUNCH(p, (k_pair_tile<M, U, W>), grid, p->tile_threads, p->tile_smem[M == 0 ? 0 : M], st, Sy, seg, dseg, out, pot_part, globalmul)
#define CHB_TILE_LAUNCH(M, U) do { if (law0) CHB_TILE_LAUNCH2(M, U, true); else CHB_TILE_LAUNCH2(M, U, false); } while (0)
#ifndef CHB_HOSTSIM
        static const int tile_var = [] { const char *e = getenv("CHB_TILE_VAR"); return e ? atoi(e) : 0; }();
        if (mode == 1 && tu && law0 && tile_var > 0) {
            const size_t sm1 = p->tile_smem[1];
            if (tile_var == 1) CHB_LAUNCH(p, (k_pair_tile_var<2, 2>), grid, p->tile_threads, sm1, st, Sy, seg, dseg, out, pot_part, globalmul);
            else if (tile_var == 2) CHB_LAUNCH(p, (k_pair_tile_var<8, 1>), grid, p->tile_threads, sm1, st, Sy, seg, dseg, out, pot_part, globalmul);
            else CHB_LAUNCH(p, (k_pair_tile_var<4, 1>), grid, p->tile_threads, sm1, st, Sy, seg, dseg, out, pot_part, globalmul);
            return CHB_OK;
        }
#endif
        if (mode == 0) { if (tu) CHB_TILE_LAUNCH(0, true); else CHB_TILE_LAUNCH(0, false); }
        else if (mode == 1) { if (tu) CHB_TILE_LAUNCH(1, true); else CHB_TILE_LAUNCH(1, false); }
        else { if (tu) CHB_TILE_LAUNCH(2, true); else CHB_TILE_LAUNCH(2, false); }
#undef CHB_TILE_LAUNCH2
#undef CHB_TILE_LAUNCH
        return CHB_OK;
    }
    if (p->sys.pair_dense) {
        const size_t wb = uni ? (size_t)p->sys.nsegm * p->sys.pair_spad * sizeof(double) : 0;  // no weight matrix in shared memory
#define CHB_DENSE_LAUNCH2(M, U, W, SM) CHB_LAUNCH(p, (k_pair_dense<M, U, W>), grid, p->dense_threads, (SM) - wb, st, Sy, seg, dseg, out, pot_part, globalmul)
#define CHB_DENSE_LAUNCH(M, U, SM) do { if (law0) CHB_DENSE_LAUNCH2(M, U, true, SM); else CHB_DENSE_LAUNCH2(M, U, false, SM); } while (0)
        if (mode == 0) { if (uni) CHB_DENSE_LAUNCH(0, true, p->dense_smem[1]); else CHB_DENSE_LAUNCH(0, false, p->dense_smem[1]); }
        else if (mode == 1) { if (uni) CHB_DENSE_LAUNCH(1, true, p->dense_smem[1]); else CHB_DENSE_LAUNCH(1, false, p->dense_smem[1]); }
        else { if (uni) CHB_DENSE_LAUNCH(2, true, p->dense_smem[2]); else CHB_DENSE_LAUNCH(2, false, p->dense_smem[2]); }
#undef CHB_DENSE_LAUNCH2
#undef CHB_DENSE_LAUNCH
        return CHB_OK;
    }
    if (mode == 0)
        CHB_LAUNCH(p, k_pair<0>, grid, p->pair_threads, p->pair_smem[0], st, Sy, seg, dseg, out, pot_part, globalmul);
    else if (mode == 1)
        CHB_LAUNCH(p, k_pair<1>, grid, p->pair_threads, p->pair_smem[1], st, Sy, seg, dseg, out, pot_part, globalmul);
    else
        CHB_LAUNCH(p, k_pair<2>, grid, p->pair_threads, p->pair_smem[2], st, Sy, seg, dseg, out, pot_part, globalmul);
    return CHB_OK;
}
static int launch_finish(chb_plan *p, int64_t nb, double *out, const double *kin_part, int nk, const double *pot_part,
                         int np, chb_stream_t st, double psign = -1.0) {
    CHB_LAUNCH(p, k_finish, (nb + 127) / 128, 128, 0, st, out, kin_part, nk, pot_part, np, (long long)nb, psign);
    return CHB_OK;
}
static int launch_kin(chb_plan *p, int64_t nb, const double *params, double *kin_out, int stride, chb_stream_t st) {
    CHB_LAUNCH(p, k_kin, nb, 256, sizeof(double) * 256, st, p->sys, params, kin_out, stride);
    return CHB_OK;
}

template <int NSEG>
static int launch_fused_n(chb_plan *p, int pmode, int64_t nb, const ChfArgs &A, chb_stream_t st) {
    switch (pmode) {
        case 0: CHB_LAUNCH(p, (k_fused<NSEG, 0>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A); break;
        case 1: CHB_LAUNCH(p, (k_fused<NSEG, 1>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A); break;
        case 2: CHB_LAUNCH(p, (k_fused<NSEG, 2>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A); break;
        default: CHB_LAUNCH(p, (k_fused<NSEG, 3>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A); break;
    }
    return CHB_OK;
}

template <int NSEG>
static int launch_fused5_n(chb_plan *p, int pmode, int64_t nb, const ChfArgs &A, chb_stream_t st) {
    switch (pmode) {
        case 0: CHB_LAUNCH(p, (k_fused5<NSEG, 0>), nb, p->fused_threads, p->fused_smem, st, p->f5, A); break;
        case 1: CHB_LAUNCH(p, (k_fused5<NSEG, 1>), nb, p->fused_threads, p->fused_smem, st, p->f5, A); break;
        case 2: CHB_LAUNCH(p, (k_fused5<NSEG, 2>), nb, p->fused_threads, p->fused_smem, st, p->f5, A); break;
        default: CHB_LAUNCH(p, (k_fused5<NSEG, 3>), nb, p->fused_threads, p->fused_smem, st, p->f5, A); break;
    }
    return CHB_OK;
}
static int launch_fused5(chb_plan *p, int pmode, int64_t nb, const ChfArgs &A0, chb_stream_t st) {
    ChfArgs A = A0;
    A.nstates = nb;
    nb = std::min<int64_t>(nb, (int64_t)p->fused_resident);  // persistent grid = one resident wave
    switch (p->f5.nseg) {
        case 1: return launch_fused5_n<1>(p, pmode, nb, A, st);
        case 2: return launch_fused5_n<2>(p, pmode, nb, A, st);
        case 3: return launch_fused5_n<3>(p, pmode, nb, A, st);
        case 4: return launch_fused5_n<4>(p, pmode, nb, A, st);
        case 5: return launch_fused5_n<5>(p, pmode, nb, A, st);
        case 6: return launch_fused5_n<6>(p, pmode, nb, A, st);
        case 7: return launch_fused5_n<7>(p, pmode, nb, A, st);
        default: return launch_fused5_n<8>(p, pmode, nb, A, st);
    }
}
// pmode: 0 energy, 1 energy + gradient, 2 Hessian-vector, 3 segment positions only
static int launch_fused(chb_plan *p, int pmode, int64_t nb, const ChfArgs &A0, chb_stream_t st) {
    if (p->fused5_active) return launch_fused5(p, pmode, nb, A0, st);
    ChfArgs A = A0;
    A.nstates = nb;
    p->fsys.ver = p->fused_enabled;
    const int64_t nb_states = nb;
    if (p->fused_enabled >= 3) nb = std::min<int64_t>(nb, (int64_t)p->fused_resident);  // persistent grid = one resident wave
    p->fsys.pf_stride = (int)std::min<int64_t>(nb, (int64_t)p->fused_resident);
    (void)nb_states;
    {
        // profiling only (phases are skipped, results are wrong): read ONCE per process, never per launch
        static const int dbg_once = [] { const char *e = getenv("CHB_FUSED_DBG"); return e ? atoi(e) : 0; }();
        p->fsys.dbg = dbg_once;
    }
    // version 4: U1 (action + gradient) and U2 (Hessian-vector on cached segment positions) with the kinetic rows = the
    // transformed rows, 16-byte aligned vectors
    if (p->fused_enabled >= 4 && p->fused4_ok && (pmode == 1 || pmode == 2) && A.vec_in && A.kin_src == A.vec_in && A.vec_out &&
        (pmode == 1 || A.seg_in) && (((uintptr_t)A.vec_in | (uintptr_t)A.vec_out | (uintptr_t)A.seg_in | (uintptr_t)A.seg_out) & 15) == 0) {
        p->fsys.ver = 4;
        {
            static const int stagger_once = [] { const char *e = getenv("CHB_FUSED_STAGGER_NS"); return e ? atoi(e) : 0; }();
            p->fsys.stagger_ns = stagger_once;  // measured: no gain (profiles/README.md), kept as a knob
        }
        if (p->fsys.nseg == 3) {
            if (pmode == 1) CHB_LAUNCH(p, (k_fused4<3, 1>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A);
            else if (A.skip || A.live_list) CHB_LAUNCH(p, (k_fused4<3, 2, 1>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A);
            else CHB_LAUNCH(p, (k_fused4<3, 2>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A);
            return CHB_OK;
        }
        if (p->fsys.nseg == 2) {
            if (pmode == 1) CHB_LAUNCH(p, (k_fused4<2, 1>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A);
            else if (A.skip || A.live_list) CHB_LAUNCH(p, (k_fused4<2, 2, 1>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A);
            else CHB_LAUNCH(p, (k_fused4<2, 2>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A);
            return CHB_OK;
        }
        if (p->fsys.nseg == 1) {
            if (pmode == 1) CHB_LAUNCH(p, (k_fused4<1, 1>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A);
            else if (A.skip || A.live_list) CHB_LAUNCH(p, (k_fused4<1, 2, 1>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A);
            else CHB_LAUNCH(p, (k_fused4<1, 2>), nb, p->fused_threads, p->fused_smem, st, p->fsys, A);
            return CHB_OK;
        }
    }
    if (p->fsys.ver > 3) p->fsys.ver = 3;
    switch (p->fsys.nseg) {
        case 1: return launch_fused_n<1>(p, pmode, nb, A, st);
        case 2: return launch_fused_n<2>(p, pmode, nb, A, st);
        case 3: return launch_fused_n<3>(p, pmode, nb, A, st);
        case 4: return launch_fused_n<4>(p, pmode, nb, A, st);
        case 5: return launch_fused_n<5>(p, pmode, nb, A, st);
        default: return launch_fused_n<6>(p, pmode, nb, A, st);
    }
}
static bool use_fused(const chb_plan *p) { return p->fused_enabled && p->fused_active; }
static ChfArgs chf_args(const double *vec_in, const double *kin_src, const double *seg_in, double *seg_out,
                        double *vec_out, double *scalar_out, int kin_from) {
    ChfArgs A;
    A.nstates = 0;  // set by launch_fused
    A.vec_in = vec_in; A.kin_src = kin_src; A.seg_in = seg_in; A.seg_out = seg_out; A.vec_out = vec_out;
    A.scalar_out = scalar_out; A.kin_from = kin_from;
    A.skip = nullptr;
    A.live_list = nullptr; A.live_count = nullptr;
    return A;
}
#define CHB_TRY(x)          \
    do {                    \
        int rc_ = (x);      \
        if (rc_) return rc_; \
    } while (0)

extern "C" int chb_params_to_segmpos_dev(chb_plan *p, int64_t batch, const double *params, double *segmpos, void *ws,
                                         size_t ws_bytes, void *stream) {
    CHB_OP_PROLOGUE("chb_params_to_segmpos_dev");
    CHB_REQUIRE(params && segmpos, CHB_ERR_INVALID_ARGUMENT, "chb_params_to_segmpos_dev: NULL array");
    CHB_SET_DEVICE();
    if (use_fused(p)) return launch_fused(p, 3, batch, chf_args(params, nullptr, nullptr, segmpos, nullptr, nullptr, 0), st);
    for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
        p->stage_c0 = c0;
        int64_t nb = std::min(chunk, batch - c0);
        CHB_TRY(launch_fwd(p, nb, params + c0 * S.nparams, segmpos + c0 * S.segsz, W, false, mask, st));
    }
    return CHB_OK;
}

static int segm_to_params_common(chb_plan *p, int64_t batch, const double *seg, double *params, void *ws,
                                 size_t ws_bytes, void *stream, int mode, const char *name) {
    CHB_OP_PROLOGUE("chb_segmpos_to_params(_T)_dev");
    (void)name;
    CHB_REQUIRE(seg && params, CHB_ERR_INVALID_ARGUMENT, "chb_segmpos_to_params(_T)_dev: NULL array");
    CHB_SET_DEVICE();
    for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
        int64_t nb = std::min(chunk, batch - c0);
        CHB_TRY(launch_adj(p, nb, seg + c0 * S.segsz, params + c0 * S.nparams, nullptr, W, mode, mask, st));
    }
    return CHB_OK;
}
extern "C" int chb_segmpos_to_params_dev(chb_plan *p, int64_t batch, const double *segmpos, double *params, void *ws,
                                         size_t ws_bytes, void *stream) {
    return segm_to_params_common(p, batch, segmpos, params, ws, ws_bytes, stream, 1, "chb_segmpos_to_params_dev");
}
extern "C" int chb_segmpos_to_params_T_dev(chb_plan *p, int64_t batch, const double *segmpos, double *params, void *ws,
                                           size_t ws_bytes, void *stream) {
    return segm_to_params_common(p, batch, segmpos, params, ws, ws_bytes, stream, 0, "chb_segmpos_to_params_T_dev");
}

extern "C" int chb_params_to_kin_nrg_dev(chb_plan *p, int64_t batch, const double *params, double *kin, void *ws,
                                         size_t ws_bytes, void *stream) {
    CHB_OP_PROLOGUE("chb_params_to_kin_nrg_dev");
    CHB_REQUIRE(params && kin, CHB_ERR_INVALID_ARGUMENT, "chb_params_to_kin_nrg_dev: NULL array");
    CHB_SET_DEVICE();
    CHB_TRY(launch_kin(p, batch, params, kin, 1, st));
    return CHB_OK;
}

extern "C" int chb_segmpos_to_pot_nrg_dev(chb_plan *p, int64_t batch, const double *segmpos, double *pot, void *ws,
                                          size_t ws_bytes, void *stream) {
    CHB_OP_PROLOGUE("chb_segmpos_to_pot_nrg_dev");
    CHB_REQUIRE(segmpos && pot, CHB_ERR_INVALID_ARGUMENT, "chb_segmpos_to_pot_nrg_dev: NULL array");
    CHB_SET_DEVICE();
    for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
        int64_t nb = std::min(chunk, batch - c0);
        CHB_TRY(launch_pair(p, 0, nb, segmpos + c0 * S.segsz, nullptr, nullptr, W.pot_part, 1.0, st));
        CHB_TRY(launch_finish(p, nb, pot + c0, nullptr, 0, W.pot_part, S.pair_ntiles, st, +1.0));
    }
    return CHB_OK;
}

// segmpos_to_hash, :2542-2560 / cdef segm_pos_to_hash :6271-6369, one exponent per call:
// hash[b] = sum_bins BinProdChargeSum/segm_size * sum_i w_i |R x_src - x_tgt|^(2 exponent)  (same loop as the potential)
extern "C" int chb_segmpos_to_hash_dev(chb_plan *p, int64_t batch, const double *segmpos, double exponent, double *hash,
                                       void *ws, size_t ws_bytes, void *stream) {
    CHB_OP_PROLOGUE("chb_segmpos_to_hash_dev");
    CHB_REQUIRE(segmpos && hash, CHB_ERR_INVALID_ARGUMENT, "chb_segmpos_to_hash_dev: NULL array");
    CHB_SET_DEVICE();
    ChbSys Sh = p->sys;  // V(x^2) = -law[3] (x^2)^(law[2] + 2) in chb_inter_law: (x^2)^exponent
    Sh.law_kind = 1;
    Sh.law[0] = 0.0; Sh.law[1] = 0.0; Sh.law[2] = exponent - 2.0; Sh.law[3] = -1.0;
    for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
        int64_t nb = std::min(chunk, batch - c0);
        CHB_TRY(launch_pair_sys(p, Sh, 0, nb, segmpos + c0 * S.segsz, nullptr, nullptr, W.pot_part, 1.0, st));
        CHB_TRY(launch_finish(p, nb, hash + c0, nullptr, 0, W.pot_part, S.pair_ntiles, st, +1.0));
    }
    return CHB_OK;
}

// One Arnoldi step of the batched GMRES for every state of the batch (chb_krylov.cuh); V is [>= j+2][batch][n], w = A v_j
// ([batch][n]), skip (nullable, one byte per state) marks states whose GMRES has stopped, h is [batch][ldh]: entries 0..j
// receive the Hessenberg column of A (the +1 of the shift included), entry j+1 the norm; V[j+1] receives the new basis vector.
static int arnoldi_launch(chb_plan *p, const char *name, int64_t batch, int64_t n, int j, int ldh, double *V, const double *w,
                          const unsigned char *skip, double *h, const ChkGivens &G, void *stream) {
    (void)name;
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, "chb_arnoldi_step_dev / chb_gmres_step_dev: NULL plan");
    CHB_REQUIRE(batch >= 0 && n >= 1 && j >= 0 && j < 62, CHB_ERR_INVALID_ARGUMENT,
                "chb_arnoldi_step_dev / chb_gmres_step_dev: need batch >= 0, n >= 1, 0 <= j < 62");
    if (batch == 0) return CHB_OK;
    CHB_REQUIRE(V && w, CHB_ERR_INVALID_ARGUMENT, "chb_arnoldi_step_dev / chb_gmres_step_dev: NULL array");
    CHB_SET_DEVICE();
    chb_stream_t st = (chb_stream_t)stream;
#ifndef CHB_HOSTSIM
    const long long ne = (n + CHK_THREADS - 1) / CHK_THREADS;
    {
        // rows through a shared-memory ring of bulk copies when they are 16-byte multiples and at least two fit
        // (CHB_ARNOLDI_RING=0: register-prefetch kernel, for A/B measurements)
        static const bool ring_on = [] { const char *e = getenv("CHB_ARNOLDI_RING"); return !(e && atoi(e) == 0); }();
        const size_t rowbytes = sizeof(double) * (size_t)n;
        const int ne_t = ne <= 4 ? 4 : ne <= 8 ? 8 : ne <= 12 ? 12 : 16;
        const size_t stagebytes = sizeof(double) * CHK_THREADS * (size_t)ne_t;  // padded row, see k_arnoldi_ring
        const size_t fit = ((size_t)216 * 1024) / stagebytes;
        const int nstage = (int)std::min<size_t>(std::min<size_t>(CHK_RING_MAX, (size_t)j + 1), fit);
        const bool aligned = (n % 2 == 0) && ((((uintptr_t)V) & 15) == 0);
        if (ring_on && aligned && ne <= 16 && j >= 1 && nstage >= 2) {
            const size_t smem = stagebytes * (size_t)nstage;
            if (ne <= 4)
                CHB_LAUNCH(p, k_arnoldi_ring<4>, batch, CHK_THREADS, smem, st, (long long)batch, (long long)n, j, ldh, V, w, skip, h, G, nstage, (unsigned)rowbytes);
            else if (ne <= 8)
                CHB_LAUNCH(p, k_arnoldi_ring<8>, batch, CHK_THREADS, smem, st, (long long)batch, (long long)n, j, ldh, V, w, skip, h, G, nstage, (unsigned)rowbytes);
            else if (ne <= 12)
                CHB_LAUNCH(p, k_arnoldi_ring<12>, batch, CHK_THREADS, smem, st, (long long)batch, (long long)n, j, ldh, V, w, skip, h, G, nstage, (unsigned)rowbytes);
            else
                CHB_LAUNCH(p, k_arnoldi_ring<16>, batch, CHK_THREADS, smem, st, (long long)batch, (long long)n, j, ldh, V, w, skip, h, G, nstage, (unsigned)rowbytes);
            return CHB_OK;
        }
    }
    if (ne <= 4)
        CHB_LAUNCH(p, k_arnoldi_reg<4>, batch, CHK_THREADS, 0, st, (long long)batch, (long long)n, j, ldh, V, w, skip, h, G);
    else if (ne <= 8)
        CHB_LAUNCH(p, k_arnoldi_reg<8>, batch, CHK_THREADS, 0, st, (long long)batch, (long long)n, j, ldh, V, w, skip, h, G);
    else if (ne <= 12)
        CHB_LAUNCH(p, k_arnoldi_reg<12>, batch, CHK_THREADS, 0, st, (long long)batch, (long long)n, j, ldh, V, w, skip, h, G);
    else if (ne <= 16)
        CHB_LAUNCH(p, k_arnoldi_reg<16>, batch, CHK_THREADS, 0, st, (long long)batch, (long long)n, j, ldh, V, w, skip, h, G);
    else {
        const size_t smem = sizeof(double) * (size_t)n;
        static const bool cluster_on = [] { const char *e = getenv("CHB_ARNOLDI_CLUSTER"); return !(e && atoi(e) == 0); }();  // A/B knob
        const long long len = ((n + CHK_CL - 1) / CHK_CL + 1) & ~1LL;  // elements per CTA of a cluster (even: 16-byte slices)
        if (smem <= 200 * 1024)
            CHB_LAUNCH(p, k_arnoldi_smem<false>, batch, CHK_THREADS, smem, st, (long long)batch, (long long)n, j, ldh, V, w, skip, h, G);
        else if (cluster_on && sizeof(double) * (size_t)len <= 200 * 1024)  // candidate vector distributed over a cluster's shared memory
            CHB_LAUNCH(p, k_arnoldi_cluster, batch * CHK_CL, CHK_THREADS, sizeof(double) * (size_t)len, st, (long long)batch, (long long)n, len, j,
                       ldh, V, w, skip, h, G);
        else  // longer still: candidate vector in its destination row
            CHB_LAUNCH(p, k_arnoldi_smem<true>, batch, CHK_THREADS, 0, st, (long long)batch, (long long)n, j, ldh, V, w, skip, h, G);
    }
#else
    (void)st;
    chk_arnoldi_host(batch, n, j, ldh, V, w, skip, h, G);
    p->launches++;
#endif
    return CHB_OK;
}
extern "C" int chb_arnoldi_step_dev(chb_plan *p, int64_t batch, int64_t n, int j, int ldh, double *V, const double *w,
                                    const unsigned char *skip, double *h, void *stream) {
    CHB_REQUIRE(h && ldh >= j + 2, CHB_ERR_INVALID_ARGUMENT, "chb_arnoldi_step_dev: h missing or ldh < j + 2");
    ChkGivens G = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, nullptr, nullptr, 0};
    return arnoldi_launch(p, "chb_arnoldi_step_dev", batch, n, j, ldh, V, w, skip, h, G, stream);
}
// One whole inner iteration of the batched GMRES(m): the Arnoldi step above, then per state the Givens rotations on the new
// Hessenberg column, the update of R and g, the iteration count and the stopping test |g[j+1]| <= target (done is both the
// skip mask on entry and the stop flag on exit).
extern "C" int chb_gmres_step_dev(chb_plan *p, int64_t batch, int64_t n, int j, int m, double *V, const double *w,
                                  unsigned char *done, double *R, double *g, double *cs, const double *target,
                                  int64_t *kdone, int *live_list_next, int *live_count_next, void *stream) {
    CHB_REQUIRE(done && R && g && cs && target && kdone && j < m, CHB_ERR_INVALID_ARGUMENT,
                "chb_gmres_step_dev: NULL array or j >= m");
    static const bool split_on = [] { const char *e = getenv("CHB_GMRES_TAIL_SPLIT"); return !(e && atoi(e) == 0); }();
    // split tail (one thread per state in a second launch) pays at large batches; with few live states the launch itself is the cost
    ChkGivens G = {R, g, cs, target, (long long *)kdone, done, m, live_list_next, live_count_next, (split_on && p && p->gmres_split_hint) ? 1 : 0};
    CHB_TRY(arnoldi_launch(p, "chb_gmres_step_dev", batch, n, j, 0, V, w, done, nullptr, G, stream));
#ifndef CHB_HOSTSIM
    if (G.split && batch > 0) CHB_LAUNCH(p, k_givens_tail, (batch + 127) / 128, 128, 0, (chb_stream_t)stream, (long long)batch, j, G);
#endif
    return CHB_OK;
}

extern "C" int chb_segmpos_to_pot_nrg_grad_dev(chb_plan *p, int64_t batch, const double *segmpos, double globalmul,
                                               double *grad, void *ws, size_t ws_bytes, void *stream) {
    CHB_OP_PROLOGUE("chb_segmpos_to_pot_nrg_grad_dev");
    CHB_REQUIRE(segmpos && grad, CHB_ERR_INVALID_ARGUMENT, "chb_segmpos_to_pot_nrg_grad_dev: NULL array");
    CHB_SET_DEVICE();
    for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
        int64_t nb = std::min(chunk, batch - c0);
        CHB_TRY(launch_pair(p, 1, nb, segmpos + c0 * S.segsz, nullptr, grad + c0 * S.segsz, nullptr, globalmul, st));
    }
    return CHB_OK;
}

extern "C" int chb_segmpos_to_pot_nrg_hess_dev(chb_plan *p, int64_t batch, const double *segmpos, const double *dsegmpos,
                                               double globalmul, double *hess, void *ws, size_t ws_bytes, void *stream) {
    CHB_OP_PROLOGUE("chb_segmpos_to_pot_nrg_hess_dev");
    CHB_REQUIRE(segmpos && dsegmpos && hess, CHB_ERR_INVALID_ARGUMENT, "chb_segmpos_to_pot_nrg_hess_dev: NULL array");
    CHB_SET_DEVICE();
    for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
        int64_t nb = std::min(chunk, batch - c0);
        CHB_TRY(launch_pair(p, 2, nb, segmpos + c0 * S.segsz, dsegmpos + c0 * S.segsz, hess + c0 * S.segsz, nullptr,
                            globalmul, st));
    }
    return CHB_OK;
}

extern "C" int chb_params_to_action_dev(chb_plan *p, int64_t batch, const double *params, double *action, void *ws,
                                        size_t ws_bytes, void *stream) {
    CHB_OP_PROLOGUE("chb_params_to_action_dev");
    CHB_REQUIRE(params && action, CHB_ERR_INVALID_ARGUMENT, "chb_params_to_action_dev: NULL array");
    CHB_SET_DEVICE();
    if (use_fused(p)) return launch_fused(p, 0, batch, chf_args(params, nullptr, nullptr, nullptr, nullptr, action, 1), st);
    for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
        int64_t nb = std::min(chunk, batch - c0);
        CHB_TRY(launch_fwd(p, nb, params + c0 * S.nparams, W.segA, W, true, mask, st));
        CHB_TRY(launch_pair(p, 0, nb, W.segA, nullptr, nullptr, W.pot_part, 1.0, st));
        CHB_TRY(launch_finish(p, nb, action + c0, W.kin_part, S.nloop, W.pot_part, S.pair_ntiles, st));
    }
    return CHB_OK;
}

extern "C" int chb_params_to_action_grad_dev(chb_plan *p, int64_t batch, const double *params, double *action,
                                             double *grad, double *segmpos_out, void *ws, size_t ws_bytes, void *stream) {
    CHB_OP_PROLOGUE("chb_params_to_action_grad_dev");
    CHB_REQUIRE(params && grad, CHB_ERR_INVALID_ARGUMENT, "chb_params_to_action_grad_dev: NULL array");
    CHB_SET_DEVICE();
    if (use_fused(p))
        return launch_fused(p, 1, batch, chf_args(params, params, nullptr, segmpos_out, grad, action, action ? 1 : 0), st);
    for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
        p->stage_c0 = c0;
        int64_t nb = std::min(chunk, batch - c0);
        double *seg = segmpos_out ? segmpos_out + c0 * S.segsz : W.segA;
        CHB_TRY(launch_fwd(p, nb, params + c0 * S.nparams, seg, W, action != nullptr, mask, st));
        CHB_TRY(launch_pair(p, 1, nb, seg, nullptr, W.segB, action ? W.pot_part : nullptr, -1.0, st));
        CHB_TRY(launch_adj(p, nb, W.segB, grad + c0 * S.nparams, params + c0 * S.nparams, W, 0, mask, st));
        if (action) CHB_TRY(launch_finish(p, nb, action + c0, W.kin_part, S.nloop, W.pot_part, S.pair_ntiles, st));
    }
    return CHB_OK;
}

extern "C" int chb_params_to_action_hess_dev(chb_plan *p, int64_t batch, const double *params, const double *dparams,
                                             double *hess, void *ws, size_t ws_bytes, void *stream) {
    CHB_OP_PROLOGUE("chb_params_to_action_hess_dev");
    CHB_REQUIRE(params && dparams && hess, CHB_ERR_INVALID_ARGUMENT, "chb_params_to_action_hess_dev: NULL array");
    CHB_SET_DEVICE();
    if (use_fused(p)) {  // two forward transforms do not fit one CTA: segment positions go through the workspace
        for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
            int64_t nb = std::min(chunk, batch - c0);
            CHB_TRY(launch_fused(p, 3, nb, chf_args(params + c0 * S.nparams, nullptr, nullptr, W.segA, nullptr, nullptr, 0), st));
            CHB_TRY(launch_fused(p, 2, nb, chf_args(dparams + c0 * S.nparams, dparams + c0 * S.nparams, W.segA, nullptr,
                                                    hess + c0 * S.nparams, nullptr, 0), st));
        }
        return CHB_OK;
    }
    for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
        int64_t nb = std::min(chunk, batch - c0);
        CHB_TRY(launch_fwd(p, nb, params + c0 * S.nparams, W.segA, W, false, mask, st));
        CHB_TRY(launch_fwd(p, nb, dparams + c0 * S.nparams, W.segB, W, false, mask, st));
        CHB_TRY(launch_pair(p, 2, nb, W.segA, W.segB, W.segC, nullptr, -1.0, st));
        CHB_TRY(launch_adj(p, nb, W.segC, hess + c0 * S.nparams, dparams + c0 * S.nparams, W, 0, mask, st));
    }
    return CHB_OK;
}

extern "C" int chb_segmpos_params_to_action_dev(chb_plan *p, int64_t batch, const double *segmpos, const double *params,
                                                double *action, void *ws, size_t ws_bytes, void *stream) {
    CHB_OP_PROLOGUE("chb_segmpos_params_to_action_dev");
    CHB_REQUIRE(segmpos && params && action, CHB_ERR_INVALID_ARGUMENT, "chb_segmpos_params_to_action_dev: NULL array");
    CHB_SET_DEVICE();
    for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
        int64_t nb = std::min(chunk, batch - c0);
        CHB_TRY(launch_kin(p, nb, params + c0 * S.nparams, W.kin_part, 1, st));
        CHB_TRY(launch_pair(p, 0, nb, segmpos + c0 * S.segsz, nullptr, nullptr, W.pot_part, 1.0, st));
        CHB_TRY(launch_finish(p, nb, action + c0, W.kin_part, 1, W.pot_part, S.pair_ntiles, st));
    }
    return CHB_OK;
}

extern "C" int chb_segmpos_params_to_action_grad_dev(chb_plan *p, int64_t batch, const double *segmpos,
                                                     const double *params, double *action, double *grad, void *ws,
                                                     size_t ws_bytes, void *stream) {
    CHB_OP_PROLOGUE("chb_segmpos_params_to_action_grad_dev");
    CHB_REQUIRE(segmpos && params && grad, CHB_ERR_INVALID_ARGUMENT,
                "chb_segmpos_params_to_action_grad_dev: NULL array");
    CHB_SET_DEVICE();
    if (use_fused(p))
        return launch_fused(p, 1, batch, chf_args(nullptr, params, segmpos, nullptr, grad, action, action ? 2 : 0), st);
    for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
        int64_t nb = std::min(chunk, batch - c0);
        if (action) CHB_TRY(launch_kin(p, nb, params + c0 * S.nparams, W.kin_part, 1, st));
        CHB_TRY(launch_pair(p, 1, nb, segmpos + c0 * S.segsz, nullptr, W.segB, action ? W.pot_part : nullptr, -1.0, st));
        CHB_TRY(launch_adj(p, nb, W.segB, grad + c0 * S.nparams, params + c0 * S.nparams, W, 0, mask, st));
        if (action) CHB_TRY(launch_finish(p, nb, action + c0, W.kin_part, 1, W.pot_part, S.pair_ntiles, st));
    }
    return CHB_OK;
}

static int hess_common(chb_plan *p, int64_t batch, const double *segmpos, const double *dparams, const unsigned char *skip,
                       const int *live_list, const int *live_count, double *hess, void *ws, size_t ws_bytes, void *stream) {
    CHB_OP_PROLOGUE("chb_segmpos_dparams_to_action_hess(_masked)_dev");
    CHB_REQUIRE(segmpos && dparams && hess, CHB_ERR_INVALID_ARGUMENT,
                "chb_segmpos_dparams_to_action_hess(_masked)_dev: NULL array");
    CHB_SET_DEVICE();
    if (use_fused(p)) {
        ChfArgs A = chf_args(dparams, dparams, segmpos, nullptr, hess, nullptr, 0);
        A.skip = skip;
        A.live_list = live_list; A.live_count = live_count;
        return launch_fused(p, 2, batch, A, st);
    }
    struct SkipGuard {  // staged kernels: CTAs of masked states return at once; the mask never outlives this call
        chb_plan *pl;
        ~SkipGuard() { pl->stage_skip = nullptr; pl->stage_c0 = 0; }
    } guard{p};
    p->stage_skip = skip;
    for (int64_t c0 = 0; c0 < batch; c0 += chunk) {
        p->stage_c0 = c0;
        int64_t nb = std::min(chunk, batch - c0);
        CHB_TRY(launch_fwd(p, nb, dparams + c0 * S.nparams, W.segB, W, false, mask, st));
        CHB_TRY(launch_pair(p, 2, nb, segmpos + c0 * S.segsz, W.segB, W.segC, nullptr, -1.0, st));
        CHB_TRY(launch_adj(p, nb, W.segC, hess + c0 * S.nparams, dparams + c0 * S.nparams, W, 0, mask, st));
    }
    return CHB_OK;
}
extern "C" int chb_segmpos_dparams_to_action_hess_dev(chb_plan *p, int64_t batch, const double *segmpos,
                                                      const double *dparams, double *hess, void *ws, size_t ws_bytes,
                                                      void *stream) {
    return hess_common(p, batch, segmpos, dparams, nullptr, nullptr, nullptr, hess, ws, ws_bytes, stream);
}
// the same operator inside the batched GMRES: states whose skip byte is non-zero (their GMRES has stopped) MAY be left out,
// their rows of hess are then not written
extern "C" int chb_segmpos_dparams_to_action_hess_masked_dev(chb_plan *p, int64_t batch, const double *segmpos,
                                                             const double *dparams, const unsigned char *skip, double *hess,
                                                             void *ws, size_t ws_bytes, void *stream) {
    return hess_common(p, batch, segmpos, dparams, skip, nullptr, nullptr, hess, ws, ws_bytes, stream);
}
// ... and with the states to process given as a list in device memory (live_list[0 .. *live_count), written by
// chb_gmres_step_dev): the persistent CTAs share the listed states evenly whatever their positions in the batch
extern "C" int chb_segmpos_dparams_to_action_hess_listed_dev(chb_plan *p, int64_t batch, const double *segmpos,
                                                             const double *dparams, const int *live_list, const int *live_count,
                                                             double *hess, void *ws, size_t ws_bytes, void *stream) {
    CHB_REQUIRE(live_list && live_count, CHB_ERR_INVALID_ARGUMENT, "chb_segmpos_dparams_to_action_hess_listed_dev: NULL list");
    return hess_common(p, batch, segmpos, dparams, nullptr, live_list, live_count, hess, ws, ws_bytes, stream);
}


// ------------------------------------------------------------------------------------------------
// Batched Newton-Krylov driver, whole solve inside the library (chb_newton.cuh): the host sequences launches and reads counters,
// every per-state decision (termination, forcing term, Armijo / Wolfe line search, best iterate) is taken on the device.
// ------------------------------------------------------------------------------------------------
struct NkArrays {  // device arrays of one solve, carved from the caller's workspace
    double *V, *w, *F, *Ft, *Fd, *xt, *dx, *seg, *R, *g, *cs, *target, *ns, *phi_t, *fmax_t, *xn2, *dxn2;
    long long *kdone;
    unsigned char *done, *mask;
    int *lists, *counts;  // lists [2][B]; counts [m + 8]: [0..m] GMRES live counts, [m + 1] live states, [m + 2], [m + 3] pending evaluations of kind 1 / 2
    void *opws;
    size_t opws_bytes;
};
static size_t nk_ws_bytes(const chb_plan *p, int64_t B, int m) {
    const size_t n = (size_t)p->sys.nparams, ss = (size_t)p->sys.segsz, Bz = (size_t)B;
    size_t b = 0;
    b += align_up(sizeof(double) * (size_t)(m + 1) * Bz * n);      // V
    b += 6 * align_up(sizeof(double) * Bz * n);                    // w F Ft Fd xt dx
    b += align_up(sizeof(double) * Bz * ss);                       // seg
    b += align_up(sizeof(double) * Bz * (size_t)(m + 1) * m) + align_up(sizeof(double) * Bz * (m + 1)) + align_up(sizeof(double) * Bz * m * 2);
    b += align_up(sizeof(double) * Bz) * 5 + align_up(sizeof(double) * Bz * NK_STRIDE);
    b += align_up(sizeof(long long) * Bz) + 2 * align_up(Bz) + align_up(sizeof(int) * 2 * Bz) + align_up(sizeof(int) * (size_t)(m + 8));
    size_t op = 0;
    chb_workspace_bytes(p, B, &op);
    return b + align_up(op) + 512;
}
static NkArrays nk_carve(const chb_plan *p, void *base, int64_t B, int m) {
    const size_t n = (size_t)p->sys.nparams, ss = (size_t)p->sys.segsz, Bz = (size_t)B;
    NkArrays a;
    char *q = (char *)(((uintptr_t)base + 255) & ~(uintptr_t)255);
    auto take = [&](size_t bytes) { char *r = q; q += align_up(bytes); return (void *)r; };
    a.V = (double *)take(sizeof(double) * (size_t)(m + 1) * Bz * n);
    a.w = (double *)take(sizeof(double) * Bz * n);
    a.F = (double *)take(sizeof(double) * Bz * n);
    a.Ft = (double *)take(sizeof(double) * Bz * n);
    a.Fd = (double *)take(sizeof(double) * Bz * n);
    a.xt = (double *)take(sizeof(double) * Bz * n);
    a.dx = (double *)take(sizeof(double) * Bz * n);
    a.seg = (double *)take(sizeof(double) * Bz * ss);
    a.R = (double *)take(sizeof(double) * Bz * (size_t)(m + 1) * m);
    a.g = (double *)take(sizeof(double) * Bz * (m + 1));
    a.cs = (double *)take(sizeof(double) * Bz * m * 2);
    a.target = (double *)take(sizeof(double) * Bz);
    a.phi_t = (double *)take(sizeof(double) * Bz);
    a.fmax_t = (double *)take(sizeof(double) * Bz);
    a.xn2 = (double *)take(sizeof(double) * Bz);
    a.dxn2 = (double *)take(sizeof(double) * Bz);
    a.ns = (double *)take(sizeof(double) * Bz * NK_STRIDE);
    a.kdone = (long long *)take(sizeof(long long) * Bz);
    a.done = (unsigned char *)take(Bz);
    a.mask = (unsigned char *)take(Bz);
    a.lists = (int *)take(sizeof(int) * 2 * Bz);
    a.counts = (int *)take(sizeof(int) * (size_t)(m + 8));
    a.opws = (void *)q;
    a.opws_bytes = 0;
    chb_workspace_bytes(p, B, &a.opws_bytes);
    return a;
}

// ---- one thread per state ----
#ifndef CHB_HOSTSIM
#define NK_FOR_STATE(b, B) const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x; if (b < (B))
#define NK_ATOMIC_INC(ptr) atomicAdd((ptr), 1)
#else
#define NK_FOR_STATE(b, B) for (long long b = sim_block_id * sim_nthreads; b < (B) && b < (sim_block_id + 1) * sim_nthreads; ++b)
static inline int nk_host_inc(int *p) { return (*p)++; }
#define NK_ATOMIC_INC(ptr) nk_host_inc(ptr)
#endif
// first residual: norms of F(x0), eta = 1e-3, best iterate = x0
CHB_KERNEL void k_nk_init(long long B, double *ns, const double *sumsq, const double *fmax) {
    NK_FOR_STATE(b, B) {
        double *st = ns + b * NK_STRIDE;
        for (int k = 0; k < NK_STRIDE; ++k) st[k] = 0.0;
        st[NK_ETA] = 1e-3;
        st[NK_FNORM] = sqrt(sumsq[b]);
        st[NK_FMAX] = fmax[b];
        const double f = st[NK_FNORM];
        st[NK_BEST] = (f == f) ? f : 1e100;
    }
}
CHB_KERNEL void k_nk_begin(long long B, double *ns, NkOpts O, int m, double *g, double *target, long long *kdone, unsigned char *done,
                           unsigned char *mask, int *list0, int *counts) {
    NK_FOR_STATE(b, B) {
        double *st = ns + b * NK_STRIDE;
        const int live = nk_begin_state(st, O);
        done[b] = live ? 0 : 1;
        mask[b] = live ? 0 : 1;
        kdone[b] = 0;
        if (live) {
            g[b * (m + 1)] = st[NK_FNORM];               // beta
            target[b] = st[NK_TOL] * st[NK_FNORM];       // rtol * |rhs|
            list0[NK_ATOMIC_INC(&counts[0])] = (int)b;
            NK_ATOMIC_INC(&counts[m + 1]);
        }
    }
}
CHB_KERNEL void k_nk_ls_start(long long B, double *ns, NkOpts O, const double *xn2, const double *dxn2, int *counts, int m) {
    NK_FOR_STATE(b, B) {
        double *st = ns + b * NK_STRIDE;
        st[NK_XNORM2] = xn2[b];
        st[NK_DXNORM2] = dxn2[b];
        nk_ls_start(st, O);
        const int t = (int)st[NK_LS_TASK];
        if (t == 1) NK_ATOMIC_INC(&counts[m + 2]);
        if (t == 2) NK_ATOMIC_INC(&counts[m + 3]);
    }
}
// mask[b] = 0 for the states that want an evaluation of this kind
CHB_KERNEL void k_nk_ls_update(long long B, double *ns, NkOpts O, int kind, const double *phi_t, const double *fmax_t, int *counts, int m) {
    NK_FOR_STATE(b, B) {
        double *st = ns + b * NK_STRIDE;
        nk_ls_update(st, O, kind, phi_t[b], fmax_t[b]);
        const int t = (int)st[NK_LS_TASK];
        if (t == 1) NK_ATOMIC_INC(&counts[m + 2]);
        if (t == 2) NK_ATOMIC_INC(&counts[m + 3]);
    }
}
// end of the iteration: forcing terms; mask[b] = 0 where the new iterate is the best so far (current_best.update, nonlin.py:13-39)
CHB_KERNEL void k_nk_end(long long B, double *ns, NkOpts O, unsigned char *mask) {
    NK_FOR_STATE(b, B) {
        double *st = ns + b * NK_STRIDE;
        const bool live = st[NK_LIVE] != 0.0;
        nk_end_state(st, O);
        bool better = false;
        if (live && st[NK_FNORM] < st[NK_BEST]) {
            st[NK_BEST] = st[NK_FNORM];
            better = true;
        }
        mask[b] = better ? 0 : 1;
    }
}
CHB_KERNEL void k_nk_finish(long long B, const double *ns, int32_t *status, int32_t *nit, double *fmax, double *fnorm2, double *best_fnorm) {
    NK_FOR_STATE(b, B) {
        const double *st = ns + b * NK_STRIDE;
        const int s = (int)st[NK_STATUS];
        status[b] = s == 0 ? 2 : s;  // still running after maxiter iterations: status 2 (the for / else of nonlin.py:109-156)
        nit[b] = (int)st[NK_NIT];
        if (fmax) fmax[b] = st[NK_FMAX];
        if (fnorm2) fnorm2[b] = st[NK_FNORM];
        if (best_fnorm) best_fnorm[b] = st[NK_BEST];
    }
}

// ---- one CTA per state: row operations ----
// op 0: out = a / |F|          (first Krylov vector)          rows with NK_LIVE
// op 1: out = a + s_eval * d   (trial point)                  rows whose mask byte is 0
// op 2: a += s_final * d, c = e  (accept: x, F)               rows with NK_LIVE
// op 3: out = a                (best iterate)                 rows whose mask byte is 0
CHB_KERNEL void k_nk_rows(long long n, int op, const double *ns, unsigned char *mask, double *out, double *a, const double *d,
                          double *c, const double *e, int nchunk, int kind, int *zero2) {
    const long long b = CHB_BLOCK_ID / nchunk;  // grid = states x nchunk: long rows are shared by several CTAs
    const long long ck = CHB_BLOCK_ID % nchunk;
    const double *st = ns + b * NK_STRIDE;
    if (op == 4) {
        // op 1 for the states whose pending line-search evaluation is of `kind` (what k_nk_mask + op 1 did in two launches): every
        // CTA of the row derives the predicate from the state, the first one records it in the mask for the kernels that follow;
        // the first CTA of the grid also clears the two "pending evaluations" counters that k_nk_ls_update will count into (the
        // host has read them: nk_read_counts synchronises the stream before this launch)
        const bool off = ((int)st[NK_LS_TASK] != kind);
#ifndef CHB_HOSTSIM
        if (threadIdx.x == 0) {
#else
        {
#endif
            if (ck == 0) mask[b] = off ? 1 : 0;
            if (CHB_BLOCK_ID == 0 && zero2) { zero2[0] = 0; zero2[1] = 0; }
        }
        if (off) return;
    }
    if ((op == 0 || op == 2) && st[NK_LIVE] == 0.0) return;
    if ((op == 1 || op == 3) && mask[b]) return;
    const double s = op == 0 ? 1.0 / st[NK_FNORM] : ((op == 1 || op == 4) ? st[NK_S_EVAL] : st[NK_S_FINAL]);
    const size_t o = (size_t)b * n;
    const long long per = (n + nchunk - 1) / nchunk, i0 = ck * per, i1 = i0 + per < n ? i0 + per : n;
#ifndef CHB_HOSTSIM
    for (long long i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
#else
    for (long long i = i0; i < i1; ++i) {
#endif
        if (op == 0) out[o + i] = a[o + i] * s;
        else if (op == 1 || op == 4) out[o + i] = fma(s, d[o + i], a[o + i]);
        else if (op == 2) { a[o + i] = fma(s, d[o + i], a[o + i]); c[o + i] = e[o + i]; }
        else out[o + i] = a[o + i];
    }
}
// per-row sum of squares (non-finite when the row has a non-finite entry) and max |.| of the rows whose mask byte is 0 (NULL: all)
CHB_KERNEL void k_nk_row_stats(long long n, const double *a, const unsigned char *mask, double *sumsq, double *maxabs,
                               long long B2, const double *a2, double *sumsq2) {
    // grid = B2 blocks for one array, 2 * B2 for two (a2 / sumsq2: the second array shares the mask; no max for it)
    long long b = CHB_BLOCK_ID;
    if (a2 && b >= B2) {
        b -= B2;
        a = a2;
        sumsq = sumsq2;
        maxabs = (double *)0;
    }
    if (mask && mask[b]) return;
    const double *r = a + (size_t)b * n;
#ifndef CHB_HOSTSIM
    __shared__ double s1[8], s2[8];
    double p = 0.0, mx = 0.0;
    for (long long e = threadIdx.x; e < n; e += 256) {
        const double x = r[e];
        p = fma(x, x, p);
        mx = fmax(mx, fabs(x));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        p += __shfl_xor_sync(0xffffffffu, p, o);
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if ((threadIdx.x & 31) == 0) { s1[threadIdx.x >> 5] = p; s2[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double P = 0.0, M = 0.0;
        for (int i = 0; i < 8; ++i) { P += s1[i]; M = fmax(M, s2[i]); }
        sumsq[b] = P;
        if (maxabs) maxabs[b] = M;
    }
#else
    double p = 0.0, mx = 0.0;
    for (long long e = 0; e < n; ++e) { p += r[e] * r[e]; mx = std::fmax(mx, std::fabs(r[e])); }
    sumsq[b] = p;
    if (maxabs) maxabs[b] = mx;
#endif
}

static int nk_read_counts(chb_plan *p, const int *d_counts, int n, int *h, chb_stream_t st) {
#ifndef CHB_HOSTSIM
    CHB_CUDA(cudaMemcpyAsync(h, d_counts, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
    CHB_CUDA(cudaStreamSynchronize(st));
#else
    (void)st;
    memcpy(h, d_counts, sizeof(int) * n);
#endif
    (void)p;
    return CHB_OK;
}
static int nk_zero(void *d, size_t bytes, chb_stream_t st) {
#ifndef CHB_HOSTSIM
    CHB_CUDA(cudaMemsetAsync(d, 0, bytes, st));
#else
    (void)st;
    memset(d, 0, bytes);
#endif
    return CHB_OK;
}
// gradient of the action at the rows whose mask byte is 0 (all rows when mask == NULL); the staged kernels compute every row
static int nk_grad(chb_plan *p, int64_t B, const double *x, double *grad, const unsigned char *mask, NkArrays &A, chb_stream_t st) {
    if (use_fused(p)) {
        ChfArgs a = chf_args(x, x, nullptr, nullptr, grad, nullptr, 0);
        a.skip = mask;
        return launch_fused(p, 1, B, a, st);
    }
    p->stage_skip = mask;
    const int rc = chb_params_to_action_grad_dev(p, B, x, nullptr, grad, nullptr, A.opws, A.opws_bytes, (void *)st);
    p->stage_skip = nullptr;
    p->stage_c0 = 0;
    return rc;
}
static int nk_segmpos(chb_plan *p, int64_t B, const double *x, double *seg, const unsigned char *mask, NkArrays &A, chb_stream_t st) {
    if (use_fused(p)) {
        ChfArgs a = chf_args(x, nullptr, nullptr, seg, nullptr, nullptr, 0);
        a.skip = mask;
        return launch_fused(p, 3, B, a, st);
    }
    p->stage_skip = mask;
    const int rc = chb_params_to_segmpos_dev(p, B, x, seg, A.opws, A.opws_bytes, (void *)st);
    p->stage_skip = nullptr;
    p->stage_c0 = 0;
    return rc;
}

extern "C" int chb_newton_workspace_bytes(const chb_plan *p, int64_t batch, int restart, size_t *bytes) {
    CHB_REQUIRE(p && bytes && batch >= 0, CHB_ERR_INVALID_ARGUMENT, "chb_newton_workspace_bytes: bad argument");
    CHB_REQUIRE(p->ready, CHB_ERR_NOT_READY, "chb_newton_workspace_bytes: call chb_plan_set_nint first");
    CHB_REQUIRE(restart >= 1 && restart <= 62, CHB_ERR_INVALID_ARGUMENT, "chb_newton_workspace_bytes: 1 <= restart <= 62");
    *bytes = nk_ws_bytes(p, std::max<int64_t>(batch, 1), restart);
    return CHB_OK;
}

extern "C" int chb_newton_krylov_dev(chb_plan *p, int64_t batch, double *x, const chb_newton_opts *opts, int32_t *status, int32_t *nit,
                                     double *fnorm_max, double *fnorm_2, double *best_x, double *best_fnorm, int64_t *counters,
                                     void *ws, size_t ws_bytes, void *stream) {
    CHB_REQUIRE(p && opts, CHB_ERR_INVALID_ARGUMENT, "chb_newton_krylov_dev: NULL plan / options");
    CHB_REQUIRE(p->ready, CHB_ERR_NOT_READY, "chb_newton_krylov_dev: call chb_plan_set_nint first");
    CHB_REQUIRE(batch >= 0, CHB_ERR_INVALID_ARGUMENT, "chb_newton_krylov_dev: negative batch");
    if (counters) counters[0] = counters[1] = counters[2] = counters[3] = 0;
    if (batch == 0) return CHB_OK;
    CHB_REQUIRE(x && status && nit, CHB_ERR_INVALID_ARGUMENT, "chb_newton_krylov_dev: NULL array");
    CHB_REQUIRE(opts->line_search >= 0 && opts->line_search <= 2, CHB_ERR_INVALID_ARGUMENT,
                "chb_newton_krylov_dev: line_search must be 0 (none), 1 (armijo) or 2 (wolfe)");
    CHB_REQUIRE(opts->tol_norm == 0 || opts->tol_norm == 1, CHB_ERR_INVALID_ARGUMENT, "chb_newton_krylov_dev: tol_norm must be 0 (max) or 1 (2-norm)");
    CHB_REQUIRE(opts->restart >= 1 && opts->restart <= 62 && opts->maxiter >= 0, CHB_ERR_INVALID_ARGUMENT,
                "chb_newton_krylov_dev: need 1 <= restart <= 62, maxiter >= 0");
    const int m = opts->restart;
    const long long n = p->sys.nparams;
    CHB_REQUIRE(ws && ws_bytes >= nk_ws_bytes(p, batch, m), CHB_ERR_WORKSPACE, "chb_newton_krylov_dev: workspace missing or too small");
    CHB_SET_DEVICE();
    chb_stream_t st = (chb_stream_t)stream;
    NkArrays A = nk_carve(p, ws, batch, m);
    NkOpts O;
    O.f_tol = opts->f_tol; O.smin = opts->smin; O.maxiter = opts->maxiter; O.restart = m;
    O.line_search = opts->line_search; O.tol_norm = opts->tol_norm; O.track_best = best_x ? 1 : 0;
    const long long B = batch;
    const unsigned gs = (unsigned)((B + 127) / 128);
    const int64_t l0 = p->launches;
    int64_t n_u1 = 0, n_u2 = 0, n_it = 0;
    int hc[4];
    if (!use_fused(p)) {
        // the staged kernels compute every state of the batch, also the ones whose solve has stopped: their Krylov vectors must
        // at least be finite numbers
        CHB_TRY(nk_zero(A.V, sizeof(double) * (size_t)(m + 1) * (size_t)B * (size_t)n, st));
        CHB_TRY(nk_zero(A.w, sizeof(double) * (size_t)B * (size_t)n, st));
    }
    // F(x0), norms, per-state scalars
    CHB_TRY(nk_grad(p, B, x, A.F, nullptr, A, st));
    n_u1 += B;
    CHB_LAUNCH(p, k_nk_row_stats, B, 256, 0, st, n, (const double *)A.F, (const unsigned char *)nullptr, A.phi_t, A.fmax_t, (long long)B, (const double *)nullptr, (double *)nullptr);
    const int nkc = (int)std::max<int64_t>(1, std::min<int64_t>(64, n / 8192));  // CTAs per row of the element-wise row kernels
    struct HintGuard {  // the hint is the driver's: other callers of chb_gmres_step_dev get the default back on every exit path
        chb_plan *pl;
        ~HintGuard() { pl->gmres_split_hint = 1; }
    } hint_guard{p};
    CHB_LAUNCH(p, k_nk_init, gs, 128, 0, st, B, A.ns, A.phi_t, A.fmax_t);
    if (best_x) {
        CHB_TRY(nk_zero(A.mask, (size_t)B, st));
        CHB_LAUNCH(p, k_nk_rows, B * nkc, 256, 0, st, n, 3, A.ns, A.mask, best_x, x, (const double *)nullptr, (double *)nullptr, (const double *)nullptr, nkc, 0, (int *)nullptr);
    }
    for (int it = 0; it < opts->maxiter; ++it) {
        CHB_TRY(nk_zero(A.counts, sizeof(int) * (size_t)(m + 8), st));
        CHB_LAUNCH(p, k_nk_begin, gs, 128, 0, st, B, A.ns, O, m, A.g, A.target, A.kdone, A.done, A.mask, A.lists, A.counts);
        CHB_TRY(nk_read_counts(p, A.counts + m + 1, 1, hc, st));
        const int nlive = hc[0];
        if (nlive == 0) break;
        ++n_it;
        p->gmres_split_hint = nlive >= 512 ? 1 : 0;
        // jacobian.setup / update (:2013-2019): segment positions of the current iterate; first Krylov vector F / |F|
        CHB_TRY(nk_segmpos(p, B, x, A.seg, A.mask, A, st));
        CHB_LAUNCH(p, k_nk_rows, B * nkc, 256, 0, st, n, 0, A.ns, A.mask, A.V, A.F, (const double *)nullptr, (double *)nullptr, (const double *)nullptr, nkc, 0, (int *)nullptr);
        // GMRES(m), one cycle, zero initial guess: per inner iteration one Hessian-vector launch over the running states and one
        // step launch (Arnoldi + Givens + stopping test + the list of the next iteration)
        for (int j = 0; j < m; ++j) {
            if (j > 0 && (j & 3) == 0) {
                CHB_TRY(nk_read_counts(p, A.counts + j, 1, hc, st));
                if (hc[0] == 0) break;
            }
            CHB_TRY(hess_common(p, B, A.seg, A.V + (size_t)j * B * n, use_fused(p) ? nullptr : A.done, A.lists + (size_t)(j & 1) * B, A.counts + j, A.w, A.opws,
                                A.opws_bytes, (void *)st));
            CHB_TRY(chb_gmres_step_dev(p, B, n, j, m, A.V, A.w, A.done, A.R, A.g, A.cs, A.target, (int64_t *)A.kdone,
                                       A.lists + (size_t)((j + 1) & 1) * B, A.counts + j + 1, (void *)st));
        }
        CHB_TRY(chb_gmres_solution_dev(p, B, n, m, A.V, A.R, A.g, (const int64_t *)A.kdone, -1.0, A.dx, (void *)st));  // dx = -J^-1 F
        // line search on phi(s) = |F(x + s dx)|^2
        CHB_LAUNCH(p, k_nk_row_stats, 2 * B, 256, 0, st, n, (const double *)x, (const unsigned char *)A.mask, A.xn2, (double *)nullptr, (long long)B,
                   (const double *)A.dx, A.dxn2);
        CHB_TRY(nk_zero(A.counts + m + 2, 2 * sizeof(int), st));
        CHB_LAUNCH(p, k_nk_ls_start, gs, 128, 0, st, B, A.ns, O, A.xn2, A.dxn2, A.counts, m);
        {   // one read: the live counts of the GMRES iterations (Hessian-vector evaluations) and the pending evaluations
            std::vector<int> all((size_t)m + 4);
            CHB_TRY(nk_read_counts(p, A.counts, m + 4, all.data(), st));
            for (int j = 0; j < m; ++j) n_u2 += all[j];
            hc[0] = all[m + 2];
            hc[1] = all[m + 3];
        }
        int guard = 0;
        while ((hc[0] > 0 || hc[1] > 0) && guard++ < 400) {
            const int use_kind = hc[1] > 0 ? 2 : 1;  // probes first: a state alternates probe / stored evaluations
            double *Fout = use_kind == 1 ? A.Ft : A.Fd;
            CHB_LAUNCH(p, k_nk_rows, B * nkc, 256, 0, st, n, 4, A.ns, A.mask, A.xt, x, (const double *)A.dx, (double *)nullptr, (const double *)nullptr, nkc, use_kind,
                       A.counts + m + 2);
            CHB_TRY(nk_grad(p, B, A.xt, Fout, A.mask, A, st));
            n_u1 += use_kind == 1 ? hc[0] : hc[1];
            CHB_LAUNCH(p, k_nk_row_stats, B, 256, 0, st, n, (const double *)Fout, (const unsigned char *)A.mask, A.phi_t, A.fmax_t, (long long)B, (const double *)nullptr, (double *)nullptr);
            CHB_LAUNCH(p, k_nk_ls_update, gs, 128, 0, st, B, A.ns, O, use_kind, A.phi_t, A.fmax_t, A.counts, m);
            CHB_TRY(nk_read_counts(p, A.counts + m + 2, 2, hc, st));
        }
        // accept: x += s dx, F = F(x + s dx) (the stored evaluation), forcing terms, best iterate
        CHB_LAUNCH(p, k_nk_rows, B * nkc, 256, 0, st, n, 2, A.ns, A.mask, (double *)nullptr, x, (const double *)A.dx, A.F, (const double *)A.Ft, nkc, 0, (int *)nullptr);
        CHB_LAUNCH(p, k_nk_end, gs, 128, 0, st, B, A.ns, O, A.mask);
        if (best_x)
            CHB_LAUNCH(p, k_nk_rows, B * nkc, 256, 0, st, n, 3, A.ns, A.mask, best_x, x, (const double *)nullptr, (double *)nullptr, (const double *)nullptr, nkc, 0, (int *)nullptr);
    }
    CHB_LAUNCH(p, k_nk_finish, gs, 128, 0, st, B, A.ns, status, nit, fnorm_max, fnorm_2, best_fnorm);
    {
        if (counters) {
            counters[0] = n_u1;
            counters[1] = n_u2;
            counters[2] = p->launches - l0;
            counters[3] = n_it;
        }
    }
#ifndef CHB_HOSTSIM
    CHB_CUDA(cudaStreamSynchronize(st));
#endif
    return CHB_OK;
}

// per-row sum of squares and max |.| of a [batch][n] array in one pass (norms of the Newton driver)
extern "C" int chb_row_stats_dev(chb_plan *p, int64_t batch, int64_t n, const double *a, double *sumsq, double *maxabs,
                                 void *stream) {
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, "chb_row_stats_dev: NULL plan");
    CHB_REQUIRE(batch >= 0 && n >= 1, CHB_ERR_INVALID_ARGUMENT, "chb_row_stats_dev: need batch >= 0, n >= 1");
    if (batch == 0) return CHB_OK;
    CHB_REQUIRE(a && sumsq, CHB_ERR_INVALID_ARGUMENT, "chb_row_stats_dev: NULL array");
    CHB_SET_DEVICE();
#ifndef CHB_HOSTSIM
    CHB_LAUNCH(p, k_row_stats, batch, 256, 0, (chb_stream_t)stream, (long long)n, a, sumsq, maxabs);
#else
    (void)stream;
    chk_row_stats_host(batch, n, a, sumsq, maxabs);
    p->launches++;
#endif
    return CHB_OK;
}

// End of a GMRES cycle for every state: back substitution on the state's own kdone x kdone triangle of R, then
// sol = scale * sum_(j < kdone) y_j v_j (scipy.sparse.linalg.gmres: y = solve_triangular(R, g); x += V^T y)
extern "C" int chb_gmres_solution_dev(chb_plan *p, int64_t batch, int64_t n, int m, const double *V, const double *R,
                                      const double *g, const int64_t *kdone, double scale, double *sol, void *stream) {
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, "chb_gmres_solution_dev: NULL plan");
    CHB_REQUIRE(batch >= 0 && n >= 1 && m >= 1 && m <= 62, CHB_ERR_INVALID_ARGUMENT,
                "chb_gmres_solution_dev: need batch >= 0, n >= 1, 1 <= m <= 62");
    if (batch == 0) return CHB_OK;
    CHB_REQUIRE(V && R && g && kdone && sol, CHB_ERR_INVALID_ARGUMENT, "chb_gmres_solution_dev: NULL array");
    CHB_SET_DEVICE();
    chb_stream_t st = (chb_stream_t)stream;
#ifndef CHB_HOSTSIM
    const int nchunk = (int)std::max<int64_t>(1, std::min<int64_t>(64, n / 8192));  // long rows: several CTAs per state
    CHB_LAUNCH(p, k_gmres_solution, batch * nchunk, 256, 0, st, (long long)batch, (long long)n, m, V, R, g, (const long long *)kdone, scale, sol, nchunk);
#else
    (void)st;
    chk_solution_host(batch, n, m, V, R, g, (const long long *)kdone, scale, sol);
    p->launches++;
#endif
    return CHB_OK;
}


// ---- post-solve reductions and small diagonal operators (chb_post.cuh), one CTA per state --------------------------
CHB_KERNEL void k_hash_multi(ChbSys S, ChbHashExp H, const double *seg, double *hash) {
    CHB_CTX_DECL;
    const long long b = CHB_BLOCK_ID;
    cta_hash_multi(ctx, S, H, seg + b * S.segsz, hash + b * H.n, (double *)CHB_SMEM);
}
CHB_KERNEL void k_center_of_mass(ChbSys S, ChbBodyTab T, const double *seg, double *com) {
    CHB_CTX_DECL;
    const long long b = CHB_BLOCK_ID;
    double *part = (double *)CHB_SMEM;
    cta_center_of_mass(ctx, S, T, seg + b * S.segsz, com + b * S.G, part, part + (size_t)2 * CHB_MAX_G * CHB_NTHREADS);
}
CHB_KERNEL void k_detect_escape(ChbSys S, ChbBodyTab T, const double *seg, double fac, int *flag) {
    CHB_CTX_DECL;
    const long long b = CHB_BLOCK_ID;
    double *part = (double *)CHB_SMEM;
    double *mm = part + (size_t)2 * CHB_MAX_G * CHB_NTHREADS;
    double *box = mm + (size_t)2 * CHB_MAX_G * S.nsegm;
    cta_detect_escape(ctx, S, T, seg + b * S.segsz, fac, flag + b, part, mm, box);
}
CHB_KERNEL void k_path_stats(ChbSys S, int nint_min, const double *seg, const double *vel, double *loop_len, double *bin_min) {
    CHB_CTX_DECL;
    const long long b = CHB_BLOCK_ID;
    double *part = (double *)CHB_SMEM;
    cta_path_stats(ctx, S, nint_min, seg + b * S.segsz, vel + b * S.segsz, loop_len + b * S.nloop, bin_min + b * S.nbin, part,
                   part + CHB_NTHREADS);
}
CHB_KERNEL void k_changevar(ChbSys S, const long long *pos_off, int mode, long long n_in, long long n_out, const double *in,
                            double *out) {
    CHB_CTX_DECL;
    const long long b = CHB_BLOCK_ID;
    cta_changevar(ctx, S, pos_off, mode, in + b * n_in, out + b * n_out);
}
// out[v][s*G + d][j] = mom[v][s*G + d][j] / segment mass  (Compute_velocities_vectorized / Compute_grad_velocities_vectorized)
CHB_KERNEL void k_scale_segm(ChbSys S, long long total, long long inner, const double *mom, double *out) {
#ifndef CHB_HOSTSIM
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
#else
    for (long long e = sim_block_id * sim_nthreads; e < total && e < (sim_block_id + 1) * sim_nthreads; ++e) {
#endif
        const int s = (int)(((e / inner) / S.G) % S.nsegm);
        const ChbLoop &L = S.loops[S.segs[s].loop];
        out[e] = mom[e] * L.invmass;
    }
}

#define CHB_POST_PROLOGUE(name)                                                                   \
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, name ": NULL plan");                                 \
    CHB_REQUIRE(p->ready, CHB_ERR_NOT_READY, name ": call chb_plan_set_nint first");              \
    CHB_REQUIRE(batch >= 0, CHB_ERR_INVALID_ARGUMENT, name ": negative batch");                   \
    if (batch == 0) return CHB_OK;                                                                \
    const ChbSys &S = p->sys;                                                                     \
    chb_stream_t st = (chb_stream_t)stream;                                                       \
    (void)S; (void)st;                                                                            \
    CHB_SET_DEVICE();

extern "C" int chb_params_to_segmvel_dev(chb_plan *p, int64_t batch, const double *params, double *segmvel, void *ws,
                                         size_t ws_bytes, void *stream) {
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, "chb_params_to_segmvel_dev: NULL plan");
    CHB_REQUIRE(p->ready, CHB_ERR_NOT_READY, "chb_params_to_segmvel_dev: call chb_plan_set_nint first");
    CHB_TRY(ensure_vel_plan(p));  // before the empty-batch return: a call with batch 0 makes chb_workspace_bytes cover this operator
    return chb_params_to_segmpos_dev(p->vel_plan, batch, params, segmvel, ws, ws_bytes, stream);
}

extern "C" int chb_params_changevar_dev(chb_plan *p, int64_t batch, const double *in, int inv, int transpose, double *out,
                                        void *stream) {
    CHB_POST_PROLOGUE("chb_params_changevar_dev");
    CHB_REQUIRE(in && out, CHB_ERR_INVALID_ARGUMENT, "chb_params_changevar_dev: NULL array");
    const int mode = (inv ? 1 : 0) + (transpose ? 2 : 0);
    const bool to_pos = (mode == 0 || mode == 3);
    const long long n_in = to_pos ? S.nparams : p->nparams_incl_o, n_out = to_pos ? p->nparams_incl_o : S.nparams;
    CHB_LAUNCH(p, k_changevar, batch, 256, 0, st, S, p->d_pos_off, mode, n_in, n_out, in, out);
    return CHB_OK;
}

extern "C" int chb_segmpos_to_hash_multi_dev(chb_plan *p, int64_t batch, const double *segmpos, int nexp, const double *exponents,
                                             double *hash, void *stream) {
    CHB_POST_PROLOGUE("chb_segmpos_to_hash_multi_dev");
    CHB_REQUIRE(segmpos && exponents && hash, CHB_ERR_INVALID_ARGUMENT, "chb_segmpos_to_hash_multi_dev: NULL array");
    CHB_REQUIRE(nexp >= 1 && nexp <= CHB_MAX_HASH, CHB_ERR_UNSUPPORTED, "chb_segmpos_to_hash_multi_dev: 1 <= nexp <= 16");
    ChbHashExp H;
    H.n = nexp;
    for (int e = 0; e < CHB_MAX_HASH; ++e) H.e[e] = e < nexp ? exponents[e] : 0.0;
    CHB_LAUNCH(p, k_hash_multi, batch, CHB_POST_THREADS, sizeof(double) * CHB_MAX_HASH * CHB_POST_THREADS, st, S, H, segmpos, hash);
    return CHB_OK;
}

extern "C" int chb_center_of_mass_dev(chb_plan *p, int64_t batch, const double *segmpos, double *com, void *stream) {
    CHB_POST_PROLOGUE("chb_center_of_mass_dev");
    CHB_REQUIRE(segmpos && com, CHB_ERR_INVALID_ARGUMENT, "chb_center_of_mass_dev: NULL array");
    CHB_REQUIRE(p->body.nbody > 0, CHB_ERR_UNSUPPORTED, "chb_center_of_mass_dev: the plan was created without the body tables (bodysegm, body_spacerot, body_timerev)");
    for (size_t k = 0; k < p->T.body_timerev.size(); ++k)
        CHB_REQUIRE(p->T.body_timerev[k] > 0 || S.segm_store > S.segm_size, CHB_ERR_INVALID_ARGUMENT,
                    "chb_center_of_mass_dev: a time-reversed body segment needs segm_store = segm_size + 1");
    const size_t smem = sizeof(double) * ((size_t)2 * CHB_MAX_G * CHB_POST_THREADS + (size_t)2 * CHB_MAX_G * S.nsegm);
    CHB_LAUNCH(p, k_center_of_mass, batch, CHB_POST_THREADS, smem, st, S, p->body, segmpos, com);
    return CHB_OK;
}

extern "C" int chb_detect_escape_dev(chb_plan *p, int64_t batch, const double *segmpos, double fac, int32_t *escaped, void *stream) {
    CHB_POST_PROLOGUE("chb_detect_escape_dev");
    CHB_REQUIRE(segmpos && escaped, CHB_ERR_INVALID_ARGUMENT, "chb_detect_escape_dev: NULL array");
    CHB_REQUIRE(p->body.nbody > 0, CHB_ERR_UNSUPPORTED, "chb_detect_escape_dev: the plan was created without the body tables (bodysegm, body_spacerot, body_timerev)");
    CHB_REQUIRE(p->body.nbody <= CHB_MAX_BODY, CHB_ERR_UNSUPPORTED, "chb_detect_escape_dev: more than 64 bodies");
    const size_t smem = sizeof(double) * ((size_t)2 * CHB_MAX_G * CHB_POST_THREADS + (size_t)2 * CHB_MAX_G * S.nsegm +
                                          (size_t)(CHB_MAX_G + 1) * p->body.nbody);
    CHB_LAUNCH(p, k_detect_escape, batch, CHB_POST_THREADS, smem, st, S, p->body, segmpos, fac, (int *)escaped);
    return CHB_OK;
}

extern "C" int chb_path_stats_dev(chb_plan *p, int64_t batch, const double *segmpos, const double *segmvel, double *loop_len,
                                  double *bin_dx_min, void *stream) {
    CHB_POST_PROLOGUE("chb_path_stats_dev");
    CHB_REQUIRE(segmpos && segmvel && loop_len && bin_dx_min, CHB_ERR_INVALID_ARGUMENT, "chb_path_stats_dev: NULL array");
    const size_t smem = sizeof(double) * ((size_t)CHB_POST_THREADS + S.nsegm);
    CHB_LAUNCH(p, k_path_stats, batch, CHB_POST_THREADS, smem, st, S, p->T.nint_min, segmpos, segmvel, loop_len, bin_dx_min);
    return CHB_OK;
}

extern "C" int chb_compute_velocities_dev(chb_plan *p, int64_t nvec, int64_t grad_ndof, const double *mom, double *out, void *stream) {
    const int64_t batch = nvec;
    CHB_POST_PROLOGUE("chb_compute_velocities_dev");
    CHB_REQUIRE(grad_ndof >= 0, CHB_ERR_INVALID_ARGUMENT, "chb_compute_velocities_dev: negative grad_ndof");
    CHB_REQUIRE(mom && out, CHB_ERR_INVALID_ARGUMENT, "chb_compute_velocities_dev: NULL array");
    const long long inner = grad_ndof > 0 ? grad_ndof : 1;
    const long long total = (long long)nvec * S.nsegm * S.G * inner;
    const long long grid = std::min<long long>((total + 255) / 256, (long long)p->num_sms * 8);
#ifdef CHB_HOSTSIM
    const long long grid_sim = (total + 255) / 256;
    CHB_LAUNCH(p, k_scale_segm, grid_sim, 256, 0, st, S, total, inner, mom, out);
    (void)grid;
#else
    CHB_LAUNCH(p, k_scale_segm, grid, 256, 0, st, S, total, inner, mom, out);
#endif
    return CHB_OK;
}

// ------------------------------------------------------------------------------------------------
// memory helpers + host-pointer operators
// ------------------------------------------------------------------------------------------------
// ODE right-hand sides (chb_ode.cuh): forces and tangent forces of the reduced system at nvec phase-space points.
// dpos == NULL: forces [nvec][nsegm*G]; else grad forces [nvec][nsegm*G][ndof].
static int ode_launch(chb_plan *p, const char *name, int64_t nvec, int64_t ndof, const double *pos, const double *dpos,
                      double *out, void *stream) {
    (void)name;
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, "chb_compute_forces_dev / chb_compute_grad_forces_dev: NULL plan");
    CHB_REQUIRE(p->ready, CHB_ERR_NOT_READY, "chb_compute_forces_dev / chb_compute_grad_forces_dev: call chb_plan_set_nint first");
    CHB_REQUIRE(nvec >= 0 && ndof >= 0, CHB_ERR_INVALID_ARGUMENT,
                "chb_compute_forces_dev / chb_compute_grad_forces_dev: negative size");
    const bool grad = dpos != nullptr;
    const long long ncol = grad ? (long long)nvec * ndof : (long long)nvec;
    if (ncol == 0) return CHB_OK;
    CHB_REQUIRE(pos && out, CHB_ERR_INVALID_ARGUMENT, "chb_compute_forces_dev / chb_compute_grad_forces_dev: NULL array");
    CHB_SET_DEVICE();
    chb_stream_t st = (chb_stream_t)stream;
    const ChbSys &S = p->sys;
#ifndef CHB_HOSTSIM
    // columns per CTA: the largest of 128 / 64 / 32 whose two [entry][column] tiles fit in shared memory, three CTAs per SM
    // when the tiles are small; vectors too long for that go straight to global memory
    const int vs = S.nsegm * S.G;
    {
        // dense gather variants: identity rotations everywhere (S.pair_w), planar, enough segments to fill warps
        static const bool dense_on = [] { const char *e = getenv("CHB_ODE_DENSE"); return !(e && atoi(e) == 0); }();
        if (dense_on && S.pair_w && S.G == 2 && S.nsegm >= 8 && S.nsegm <= 256) {
            if (!grad) {
                const int nv_tile = std::max(1, 256 / S.nsegm), thr = nv_tile * S.nsegm;
                const long long ntile = (nvec + nv_tile - 1) / nv_tile;
                const long long grid = std::min<long long>(ntile, (long long)p->num_sms * (2048 / ((thr + 31) / 32 * 32)));
                CHB_LAUNCH(p, k_ode_forces_dense<2>, grid, thr, sizeof(double) * nv_tile * vs, st, S, (long long)nvec, pos, out);
                return CHB_OK;
            }
            if (ndof >= 16) {
                const long long nitem = (long long)nvec * ((ndof + CHB_ODE_KC - 1) / CHB_ODE_KC);
                const size_t sb = sizeof(double) * (size_t)vs * (CHB_ODE_KC + 1);
                const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (size_t)(200 * 1024) / (sb + 1024)));
                const long long grid = std::min<long long>(nitem, (long long)p->num_sms * per_sm);
                CHB_LAUNCH(p, k_ode_grad_forces_dense<2>, grid, 256, sb, st, S, (long long)nvec, (long long)ndof, pos, dpos, out);
                return CHB_OK;
            }
        }
    }
    int T = 128;
    auto tile_bytes = [&](int t) { return (size_t)2 * vs * (t + 1) * sizeof(double); };
    while (T > 32 && tile_bytes(T) > p->smem_limit - 1024) T >>= 1;
    static const bool no_tiles = [] { const char *e = getenv("CHB_ODE_NOSMEM"); return e && atoi(e) != 0; }();  // tests
    const bool smem = !no_tiles && tile_bytes(T) <= p->smem_limit - 1024;
    const size_t sb = smem ? tile_bytes(T) : 0;
    const long long ntile = (ncol + T - 1) / T;
    const int per_sm = smem ? (int)std::max<size_t>(1, std::min<size_t>(16, (size_t)(200 * 1024) / (sb + 1024))) : 8;
    const long long grid = std::min<long long>(ntile, (long long)p->num_sms * per_sm);
    if (grad) {
        if (smem) CHB_LAUNCH(p, (k_ode_forces<true, true>), grid, T, sb, st, S, ncol, (long long)ndof, pos, dpos, out);
        else CHB_LAUNCH(p, (k_ode_forces<true, false>), grid, T, sb, st, S, ncol, (long long)ndof, pos, dpos, out);
    } else {
        if (smem) CHB_LAUNCH(p, (k_ode_forces<false, true>), grid, T, sb, st, S, ncol, 1LL, pos, dpos, out);
        else CHB_LAUNCH(p, (k_ode_forces<false, false>), grid, T, sb, st, S, ncol, 1LL, pos, dpos, out);
    }
#else
    (void)st;
    if (grad) ode_forces_host<true>(S, ncol, ndof, pos, dpos, out);
    else ode_forces_host<false>(S, ncol, 1, pos, dpos, out);
    p->launches++;
#endif
    return CHB_OK;
}
extern "C" int chb_compute_forces_dev(chb_plan *p, int64_t nvec, const double *pos, double *forces, void *stream) {
    return ode_launch(p, "chb_compute_forces_dev", nvec, 1, pos, nullptr, forces, stream);
}
extern "C" int chb_compute_grad_forces_dev(chb_plan *p, int64_t nvec, int64_t grad_ndof, const double *pos, const double *dpos,
                                           double *grad_forces, void *stream) {
    CHB_REQUIRE(nvec * grad_ndof == 0 || dpos, CHB_ERR_INVALID_ARGUMENT, "chb_compute_grad_forces_dev: NULL dpos");
    if (nvec * grad_ndof == 0) return CHB_OK;
    return ode_launch(p, "chb_compute_grad_forces_dev", nvec, grad_ndof, pos, dpos, grad_forces, stream);
}
// Host-pointer forms: the arrays travel to the plan's device and back inside the call (dpos / grad_forces NULL: forces only).
extern "C" int chb_compute_forces_host(chb_plan *p, int64_t nvec, int64_t grad_ndof, const double *pos, const double *dpos,
                                       double *forces, double *grad_forces) {
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, "chb_compute_forces_host: NULL plan");
    CHB_REQUIRE(p->ready, CHB_ERR_NOT_READY, "chb_compute_forces_host: call chb_plan_set_nint first");
    CHB_REQUIRE(nvec >= 0 && grad_ndof >= 0, CHB_ERR_INVALID_ARGUMENT, "chb_compute_forces_host: negative size");
    CHB_REQUIRE((dpos == nullptr) == (grad_forces == nullptr), CHB_ERR_INVALID_ARGUMENT,
                "chb_compute_forces_host: dpos and grad_forces go together");
    if (nvec == 0) return CHB_OK;
    CHB_REQUIRE(pos && (forces || grad_forces), CHB_ERR_INVALID_ARGUMENT, "chb_compute_forces_host: NULL array");
    const size_t vs = (size_t)p->sys.nsegm * p->sys.G;
    const size_t nb_pos = sizeof(double) * vs * (size_t)nvec, nb_grad = dpos ? nb_pos * (size_t)grad_ndof : 0;
    void *d_pos = nullptr, *d_out = nullptr, *d_dpos = nullptr, *d_gout = nullptr;
    int rc = dev_alloc(p, nb_pos, &d_pos);
    if (!rc) rc = dev_alloc(p, nb_pos, &d_out);
    if (!rc && nb_grad) rc = dev_alloc(p, nb_grad, &d_dpos);
    if (!rc && nb_grad) rc = dev_alloc(p, nb_grad, &d_gout);
    if (!rc) rc = dev_upload(d_pos, pos, nb_pos);
    if (!rc && nb_grad) rc = dev_upload(d_dpos, dpos, nb_grad);
    if (!rc && forces) rc = chb_compute_forces_dev(p, nvec, (const double *)d_pos, (double *)d_out, nullptr);
    if (!rc && nb_grad)
        rc = chb_compute_grad_forces_dev(p, nvec, grad_ndof, (const double *)d_pos, (const double *)d_dpos, (double *)d_gout, nullptr);
    if (!rc && forces) rc = chb_memcpy_d2h(p, forces, d_out, nb_pos, nullptr);
    if (!rc && nb_grad) rc = chb_memcpy_d2h(p, grad_forces, d_gout, nb_grad, nullptr);
    if (!rc) rc = chb_stream_synchronize(p, nullptr);
    if (d_pos) dev_free(d_pos);
    if (d_out) dev_free(d_out);
    if (d_dpos) dev_free(d_dpos);
    if (d_gout) dev_free(d_gout);
    return rc;
}

extern "C" int chb_malloc_device(chb_plan *p, size_t bytes, void **out) {
    CHB_REQUIRE(p && out, CHB_ERR_INVALID_ARGUMENT, "chb_malloc_device: NULL argument");
    return dev_alloc(p, bytes, out);
}
extern "C" int chb_free_device(chb_plan *p, void *ptr) {
    (void)p;
    if (ptr) dev_free(ptr);
    return CHB_OK;
}
extern "C" int chb_malloc_pinned(size_t bytes, void **out) {
    CHB_REQUIRE(out, CHB_ERR_INVALID_ARGUMENT, "chb_malloc_pinned: NULL argument");
    if (bytes == 0) bytes = 16;
#ifndef CHB_HOSTSIM
    CHB_CUDA(cudaMallocHost(out, bytes));
#else
    *out = malloc(bytes);
#endif
    return CHB_OK;
}
extern "C" int chb_free_pinned(void *ptr) {
#ifndef CHB_HOSTSIM
    if (ptr) cudaFreeHost(ptr);
#else
    free(ptr);
#endif
    return CHB_OK;
}
extern "C" int chb_memcpy_h2d(chb_plan *p, void *dst, const void *src, size_t bytes, void *stream) {
    CHB_REQUIRE(p && (bytes == 0 || (dst && src)), CHB_ERR_INVALID_ARGUMENT, "chb_memcpy_h2d: NULL argument");
#ifndef CHB_HOSTSIM
    CHB_CUDA(cudaSetDevice(p->device));
    CHB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
#else
    (void)stream;
    memcpy(dst, src, bytes);
#endif
    return CHB_OK;
}
extern "C" int chb_memcpy_d2h(chb_plan *p, void *dst, const void *src, size_t bytes, void *stream) {
    CHB_REQUIRE(p && (bytes == 0 || (dst && src)), CHB_ERR_INVALID_ARGUMENT, "chb_memcpy_d2h: NULL argument");
#ifndef CHB_HOSTSIM
    CHB_CUDA(cudaSetDevice(p->device));
    CHB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
#else
    (void)stream;
    memcpy(dst, src, bytes);
#endif
    return CHB_OK;
}
extern "C" int chb_stream_synchronize(chb_plan *p, void *stream) {
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, "chb_stream_synchronize: NULL plan");
#ifndef CHB_HOSTSIM
    CHB_CUDA(cudaSetDevice(p->device));
    CHB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
#else
    (void)stream;
#endif
    return CHB_OK;
}

static int ensure_stage(chb_plan *p, size_t bytes) {
    if (bytes <= p->stage_dev_bytes) return CHB_OK;
    if (p->stage_dev) dev_free(p->stage_dev);
    p->stage_dev = nullptr;
    p->stage_dev_bytes = 0;
    size_t want = bytes + bytes / 4;
    int rc = dev_alloc(p, want, &p->stage_dev);
    if (rc) return rc;
    p->stage_dev_bytes = want;
    return CHB_OK;
}

// element counts per state of the arrays of a host operator
static int host_op_shapes(const chb_plan *p, int op, bool have_out0, size_t *n_in0, size_t *n_in1, size_t *n_out0,
                          size_t *n_out1) {
    const size_t np = (size_t)p->sys.nparams, ns = (size_t)p->sys.segsz;
    *n_in0 = *n_in1 = *n_out0 = *n_out1 = 0;
    switch (op) {
        case CHB_OP_PARAMS_TO_SEGMPOS:
        case CHB_OP_PARAMS_TO_SEGMVEL: *n_in0 = np; *n_out0 = ns; break;
        case CHB_OP_SEGMPOS_TO_PARAMS:
        case CHB_OP_SEGMPOS_TO_PARAMS_T: *n_in0 = ns; *n_out0 = np; break;
        case CHB_OP_ACTION: *n_in0 = np; *n_out0 = 1; break;
        case CHB_OP_ACTION_GRAD: *n_in0 = np; *n_out0 = have_out0 ? 1 : 0; *n_out1 = np; break;
        case CHB_OP_ACTION_HESS: *n_in0 = np; *n_in1 = np; *n_out0 = np; break;
        case CHB_OP_SEGMPOS_ACTION: *n_in0 = ns; *n_in1 = np; *n_out0 = 1; break;
        case CHB_OP_SEGMPOS_ACTION_GRAD: *n_in0 = ns; *n_in1 = np; *n_out0 = have_out0 ? 1 : 0; *n_out1 = np; break;
        case CHB_OP_SEGMPOS_ACTION_HESS: *n_in0 = ns; *n_in1 = np; *n_out0 = np; break;
        case CHB_OP_KIN: *n_in0 = np; *n_out0 = 1; break;
        case CHB_OP_POT: *n_in0 = ns; *n_out0 = 1; break;
        case CHB_OP_HASH: *n_in0 = ns; *n_out0 = 1; break;
        case CHB_OP_POT_GRAD: *n_in0 = ns; *n_out0 = ns; break;
        case CHB_OP_POT_HESS: *n_in0 = ns; *n_in1 = ns; *n_out0 = ns; break;
        default: return chb_fail(CHB_ERR_INVALID_ARGUMENT, "chb_run_host: unknown op");
    }
    return CHB_OK;
}

static int host_op_dispatch(chb_plan *p, int op, int64_t nb, const double *d_in0, const double *d_in1, double scalar,
                            double *d_out0, double *d_out1, bool have_out0, void *d_ws, size_t wsb, void *st) {
    switch (op) {
        case CHB_OP_PARAMS_TO_SEGMPOS: return chb_params_to_segmpos_dev(p, nb, d_in0, d_out0, d_ws, wsb, st);
        case CHB_OP_PARAMS_TO_SEGMVEL: return chb_params_to_segmvel_dev(p, nb, d_in0, d_out0, d_ws, wsb, st);
        case CHB_OP_SEGMPOS_TO_PARAMS: return chb_segmpos_to_params_dev(p, nb, d_in0, d_out0, d_ws, wsb, st);
        case CHB_OP_SEGMPOS_TO_PARAMS_T: return chb_segmpos_to_params_T_dev(p, nb, d_in0, d_out0, d_ws, wsb, st);
        case CHB_OP_ACTION: return chb_params_to_action_dev(p, nb, d_in0, d_out0, d_ws, wsb, st);
        case CHB_OP_ACTION_GRAD:
            return chb_params_to_action_grad_dev(p, nb, d_in0, have_out0 ? d_out0 : nullptr, d_out1, nullptr, d_ws, wsb, st);
        case CHB_OP_ACTION_HESS: return chb_params_to_action_hess_dev(p, nb, d_in0, d_in1, d_out0, d_ws, wsb, st);
        case CHB_OP_SEGMPOS_ACTION: return chb_segmpos_params_to_action_dev(p, nb, d_in0, d_in1, d_out0, d_ws, wsb, st);
        case CHB_OP_SEGMPOS_ACTION_GRAD:
            return chb_segmpos_params_to_action_grad_dev(p, nb, d_in0, d_in1, have_out0 ? d_out0 : nullptr, d_out1, d_ws, wsb, st);
        case CHB_OP_SEGMPOS_ACTION_HESS: return chb_segmpos_dparams_to_action_hess_dev(p, nb, d_in0, d_in1, d_out0, d_ws, wsb, st);
        case CHB_OP_KIN: return chb_params_to_kin_nrg_dev(p, nb, d_in0, d_out0, d_ws, wsb, st);
        case CHB_OP_POT: return chb_segmpos_to_pot_nrg_dev(p, nb, d_in0, d_out0, d_ws, wsb, st);
        case CHB_OP_HASH: return chb_segmpos_to_hash_dev(p, nb, d_in0, scalar, d_out0, d_ws, wsb, st);
        case CHB_OP_POT_GRAD: return chb_segmpos_to_pot_nrg_grad_dev(p, nb, d_in0, scalar, d_out0, d_ws, wsb, st);
        case CHB_OP_POT_HESS: return chb_segmpos_to_pot_nrg_hess_dev(p, nb, d_in0, d_in1, scalar, d_out0, d_ws, wsb, st);
    }
    return chb_fail(CHB_ERR_INVALID_ARGUMENT, "chb_run_host: unknown op");
}

// Host-pointer operator: the batch is cut into slices that travel through CHB_HOST_SLOTS stream slots
// (H2D, kernels, D2H of a slice are ordered on its stream; slices on different streams overlap), so that with
// pinned host buffers (chb_malloc_pinned) the PCIe link runs in both directions while the SMs compute.
extern "C" int chb_run_host(chb_plan *p, int op, int64_t batch, const double *in0, const double *in1, double scalar,
                            double *out0, double *out1) {
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, "chb_run_host: NULL plan");
    CHB_REQUIRE(p->ready, CHB_ERR_NOT_READY, "chb_run_host: call chb_plan_set_nint first");
    CHB_REQUIRE(batch >= 0, CHB_ERR_INVALID_ARGUMENT, "chb_run_host: negative batch");
    if (batch == 0) return CHB_OK;
    std::lock_guard<std::mutex> lock(p->host_mutex);
    if (op == CHB_OP_PARAMS_TO_SEGMVEL) CHB_TRY(ensure_vel_plan(p));  // before the workspace query below
    size_t n_in0, n_in1, n_out0, n_out1;
    const bool have_out0 = out0 != nullptr;
    CHB_TRY(host_op_shapes(p, op, have_out0, &n_in0, &n_in1, &n_out0, &n_out1));
    CHB_REQUIRE(in0 && (n_in1 == 0 || in1) && (n_out0 == 0 || out0) && (n_out1 == 0 || out1), CHB_ERR_INVALID_ARGUMENT,
                "chb_run_host: NULL array for this op");
    // slice size: ~48 MB of the widest array, a multiple of the SM count when the batch allows it
    const size_t widest = std::max(std::max(n_in0, n_in1), std::max(n_out0, n_out1)) * 8;
    int64_t slice = (int64_t)std::max<size_t>(1, ((size_t)48 << 20) / std::max<size_t>(widest, 1));
    if (slice >= p->num_sms) slice -= slice % p->num_sms;
    slice = std::min<int64_t>(slice, batch);
    const int nslots = (int)std::min<int64_t>(CHB_HOST_SLOTS, (batch + slice - 1) / slice);
    size_t wsb = 0;
    CHB_TRY(chb_workspace_bytes(p, slice, &wsb));
    const size_t S = (size_t)slice;
    const size_t b_in0 = align_up(S * n_in0 * 8), b_in1 = align_up(S * n_in1 * 8),
                 b_out0 = align_up(S * std::max<size_t>(n_out0, 1) * 8), b_out1 = align_up(S * n_out1 * 8);
    const size_t per_slot = b_in0 + b_in1 + b_out0 + b_out1 + align_up(wsb);
    CHB_TRY(ensure_stage(p, per_slot * nslots + 512));
#ifndef CHB_HOSTSIM
    CHB_CUDA(cudaSetDevice(p->device));
    for (int i = 0; i < nslots; ++i)
        if (!p->host_streams[i]) CHB_CUDA(cudaStreamCreateWithFlags(&p->host_streams[i], cudaStreamNonBlocking));
#endif
    char *base = (char *)(((uintptr_t)p->stage_dev + 255) & ~(uintptr_t)255);
    int rc = CHB_OK;
    int64_t c0 = 0;
    for (int it = 0; c0 < batch && rc == CHB_OK; ++it, c0 += slice) {
        const int64_t nb = std::min<int64_t>(slice, batch - c0);
        const int slot = it % nslots;
        char *q = base + per_slot * slot;
        double *d_in0 = (double *)q; q += b_in0;
        double *d_in1 = (double *)q; q += b_in1;
        double *d_out0 = (double *)q; q += b_out0;
        double *d_out1 = (double *)q; q += b_out1;
        void *d_ws = q;
#ifndef CHB_HOSTSIM
        void *st = (void *)p->host_streams[slot];
#else
        void *st = nullptr;
#endif
        const size_t B = (size_t)nb;
        rc = chb_memcpy_h2d(p, d_in0, in0 + (size_t)c0 * n_in0, B * n_in0 * 8, st);
        if (!rc && n_in1) rc = chb_memcpy_h2d(p, d_in1, in1 + (size_t)c0 * n_in1, B * n_in1 * 8, st);
        if (!rc) rc = host_op_dispatch(p, op, nb, d_in0, d_in1, scalar, d_out0, d_out1, have_out0, d_ws, wsb, st);
        if (!rc && n_out0) rc = chb_memcpy_d2h(p, out0 + (size_t)c0 * n_out0, d_out0, B * n_out0 * 8, st);
        if (!rc && n_out1) rc = chb_memcpy_d2h(p, out1 + (size_t)c0 * n_out1, d_out1, B * n_out1 * 8, st);
    }
#ifndef CHB_HOSTSIM
    for (int i = 0; i < nslots; ++i) {
        cudaError_t e = cudaStreamSynchronize(p->host_streams[i]);
        if (e != cudaSuccess && rc == CHB_OK) rc = chb_fail(CHB_ERR_CUDA, std::string("chb_run_host: ") + cudaGetErrorString(e));
    }
#endif
    return rc;
}

// ------------------------------------------------------------------------------------------------
// Hessian-vector product without any allocation by the caller or the library (no counterpart in the reference):
// the operator's scratch (segment positions of both arguments and of the result, partial sums)
// lives in ONE file-scope device array per GPU that loads the library, CHB_NOALLOC_SCRATCH_BYTES long, held until exit.
// Every call on a device shares it and finds the last call's contents there: nothing in it has to start at zero (every kernel of
// the operator writes what the next one reads), and two calls that overlap on the GPU (different streams, different plans on one
// device) race on it -- callers serialise them.
// ------------------------------------------------------------------------------------------------
#ifndef CHB_HOSTSIM
static __device__ __align__(256) unsigned char chb_noalloc_scratch[CHB_NOALLOC_SCRATCH_BYTES];
#else
alignas(256) static unsigned char chb_noalloc_scratch[CHB_NOALLOC_SCRATCH_BYTES];
#endif
#define CHB_NOALLOC_MAX_DEVICES 64
static void *g_noalloc_ptr[CHB_NOALLOC_MAX_DEVICES];  // device address of the array on each device (host code cannot use its name)
static std::mutex g_noalloc_mutex;

// set-up, once per device, with that device current: the device address of the scratch array
static int noalloc_scratch_address(chb_plan *p, void **out) {
    CHB_REQUIRE(p->device >= 0 && p->device < CHB_NOALLOC_MAX_DEVICES, CHB_ERR_UNSUPPORTED,
                "chb_params_to_action_hess_noalloc_dev: device index beyond the per-device table");
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[p->device]) {
#ifndef CHB_HOSTSIM
        CHB_CUDA(cudaSetDevice(p->device));
        void *ptr = nullptr;
        CHB_CUDA(cudaGetSymbolAddress((void **)&ptr, chb_noalloc_scratch));
        CHB_REQUIRE(ptr != nullptr, CHB_ERR_CUDA, "chb_params_to_action_hess_noalloc_dev: cudaGetSymbolAddress returned NULL");
        g_noalloc_ptr[p->device] = ptr;
#else
        g_noalloc_ptr[p->device] = (void *)chb_noalloc_scratch;
#endif
    }
    *out = g_noalloc_ptr[p->device];
    return CHB_OK;
}

extern "C" int chb_noalloc_scratch_bytes(size_t *bytes_out) {
    CHB_REQUIRE(bytes_out, CHB_ERR_INVALID_ARGUMENT, "chb_noalloc_scratch_bytes: NULL argument");
    *bytes_out = (size_t)CHB_NOALLOC_SCRATCH_BYTES;
    return CHB_OK;
}

extern "C" int chb_params_to_action_hess_noalloc_dev(chb_plan *p, int64_t batch, const double *params, const double *dparams,
                                                     double *hess, void *stream) {
    CHB_REQUIRE(p, CHB_ERR_INVALID_ARGUMENT, "chb_params_to_action_hess_noalloc_dev: NULL plan");
    CHB_REQUIRE(p->ready, CHB_ERR_NOT_READY, "chb_params_to_action_hess_noalloc_dev: call chb_plan_set_nint first");
    CHB_REQUIRE(batch >= 0, CHB_ERR_INVALID_ARGUMENT, "chb_params_to_action_hess_noalloc_dev: negative batch");
    if (batch == 0) return CHB_OK;
    CHB_REQUIRE(params && dparams && hess, CHB_ERR_INVALID_ARGUMENT, "chb_params_to_action_hess_noalloc_dev: NULL array");
    // what chb_workspace_bytes asks the caller of the workspace variant for (the operator works in chunks of states: the need
    // follows the chunk, chb_plan_set_chunk lowers it)
    const size_t need = ws_bytes_for(p, plan_chunk(p, batch)) + 256;
    if (need > (size_t)CHB_NOALLOC_SCRATCH_BYTES)
        return chb_fail(CHB_ERR_WORKSPACE, "chb_params_to_action_hess_noalloc_dev: " + std::to_string(batch) + " states of this level need " +
                                               std::to_string(need) + " bytes of scratch, the built-in array holds " +
                                               std::to_string((size_t)CHB_NOALLOC_SCRATCH_BYTES) +
                                               " (CHB_NOALLOC_SCRATCH_BYTES): lower the batch or chb_plan_set_chunk, or use "
                                               "chb_params_to_action_hess_dev with a workspace of your own");
    void *scratch = nullptr;
    CHB_TRY(noalloc_scratch_address(p, &scratch));
    return chb_params_to_action_hess_dev(p, batch, params, dparams, hess, scratch, (size_t)CHB_NOALLOC_SCRATCH_BYTES, stream);
}
