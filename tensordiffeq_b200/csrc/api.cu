// C-ABI of libtdq_b200.so (declarations and reference citations: include/tdq_b200.h).
// Host side of the hot path: owns device state, lowers terms onto the compiled
// instantiations of the fused jet kernel, sequences reduce / all-reduce / Adam / L-BFGS.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/tdq_b200.h"
#include "optim_kernels.cuh"
#include "variants.h"
#include "tc_image.cuh"
#include "vm_poly.h"

using namespace tdq;

static thread_local std::string g_last_error;

#define TDQ_FAIL(h, ...)                                   \
    do {                                                   \
        char _b[512];                                      \
        snprintf(_b, sizeof(_b), __VA_ARGS__);             \
        if (h) (h)->err = _b;                              \
        g_last_error = _b;                                 \
        return 1;                                          \
    } while (0)

#define CK(h, call)                                                                                   \
    do {                                                                                              \
        cudaError_t _e = (call);                                                                      \
        if (_e != cudaSuccess) TDQ_FAIL(h, "CUDA error %s at %s:%d (%s)", cudaGetErrorString(_e), __FILE__, __LINE__, #call); \
    } while (0)

// ---------------------------------------------------------------------------
// variant registry
// ---------------------------------------------------------------------------
namespace tdq {
static const Variant* const kTables[] = {kVariants_000, kVariants_110, kVariants_210,
                                         kVariants_220, kVariants_410, kVariants_222, kVariants_22m};
int n_variants() { return (int)(sizeof(kTables) / sizeof(kTables[0])) * kVariantsPerCfg; }
const Variant* variant_at(int i) { return &kTables[i / kVariantsPerCfg][i % kVariantsPerCfg]; }
const Variant* find_variant(int k0, int k1, int k2, int hp, int bwd, int mixed) {
    const Variant* best = nullptr;
    for (int i = 0; i < n_variants(); ++i) {
        const Variant* v = variant_at(i);
        if (v->hp != hp || v->bwd != bwd) continue;
        if ((v->k2 < 0) != (mixed != 0)) continue;          // the mixed-partial configuration carries the tag k2 = -1
        if (v->k0 < k0 || v->k1 < k1 || (!mixed && v->k2 < k2)) continue;
        if (!best || v->C < best->C) best = v;
    }
    return best;
}
}  // namespace tdq

// ---------------------------------------------------------------------------
// NCCL through dlopen (no link-time dependency; torch's bundled libnccl.so.2 is
// picked up when the host process already loaded it)
// ---------------------------------------------------------------------------
struct NcclId { char internal[128]; };
typedef int (*nccl_get_uid_t)(NcclId*);
typedef int (*nccl_init_rank_t)(void**, int, NcclId, int);
typedef int (*nccl_allreduce_t)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*nccl_allgather_t)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef int (*nccl_destroy_t)(void*);
typedef const char* (*nccl_errstr_t)(int);
struct NcclApi {
    void* lib = nullptr;
    nccl_get_uid_t get_uid = nullptr;
    nccl_init_rank_t init_rank = nullptr;
    nccl_allreduce_t allreduce = nullptr;
    nccl_allgather_t allgather = nullptr;
    nccl_destroy_t destroy = nullptr;
    nccl_errstr_t errstr = nullptr;
};
static NcclApi g_nccl;
static const char* nccl_load() {
    if (g_nccl.lib) return nullptr;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
        g_nccl.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.lib) break;
    }
    if (!g_nccl.lib) return "libnccl.so.2 not found (dlopen)";
    g_nccl.get_uid = (nccl_get_uid_t)dlsym(g_nccl.lib, "ncclGetUniqueId");
    g_nccl.init_rank = (nccl_init_rank_t)dlsym(g_nccl.lib, "ncclCommInitRank");
    g_nccl.allreduce = (nccl_allreduce_t)dlsym(g_nccl.lib, "ncclAllReduce");
    g_nccl.allgather = (nccl_allgather_t)dlsym(g_nccl.lib, "ncclAllGather");
    g_nccl.destroy = (nccl_destroy_t)dlsym(g_nccl.lib, "ncclCommDestroy");
    g_nccl.errstr = (nccl_errstr_t)dlsym(g_nccl.lib, "ncclGetErrorString");
    if (!g_nccl.get_uid || !g_nccl.init_rank || !g_nccl.allreduce) return "NCCL symbols missing";
    return nullptr;
}

// ---------------------------------------------------------------------------
// state
// ---------------------------------------------------------------------------
struct Term {
    int n_dims = 0;
    int orders[TDQ_MAX_DIMS] = {0, 0, 0};
    std::vector<tdq_op> ops;      // as given in the spec (JET channels in the spec's own numbering)
    int out_slot = 0, f_slot = -1, pair = 0, mixed = 0;
    const Variant* vb = nullptr;  // smallest compiled configuration covering this term alone (its cost weight)
    int group = -1;               // launch group evaluating this term (build_plan)
    long long n = 0, n_global = 0, cap = 0;
    long long r_start = 0, r_count = -1, r_global = 0;   // minibatch range (r_count < 0: whole set)
    float* X = nullptr;           // [d][n]
    float* X_next = nullptr;      // shadow point set being filled by tdq_prefetch_term_points (same capacity as X)
    long long next_n = -1;        // its point count; < 0: nothing staged
    float* target = nullptr;
    float* lam = nullptr;
    float* lam_m = nullptr;
    float* lam_v = nullptr;
    float* lam_grad = nullptr;
    float* f_out = nullptr;
    bool has_target = false, has_lam = false;
};

// Terms evaluated by ONE launch of the jet kernel (KParams::seg).  The big residual term takes the small BC / IC
// terms of the same loss along: they add a few tiles to its tile queue instead of costing a launch each.
struct Group {
    std::vector<int> terms;       // indices into tdq_handle::terms; the heaviest first
    const Variant* vb = nullptr;  // compiled configuration covering every term of the group (fwd + bwd)
    const TcVariant* tc = nullptr;   // tensor-core kernel for the same configuration, when the net qualifies
    struct Prog { int ops_off, n_ops, out_slot, f_slot, tc_off, tc_n_ops, tc_out, tc_f; };
    std::vector<Prog> prog;       // per term of the group
    tdq_op* ops_dev = nullptr;    // the terms' programs back to back, JET channels in vb's numbering
    TcOp* tcops_dev = nullptr;    // ... compiled for the tensor-core kernel (vm_poly.h)
    int n_ops = 0, tc_n_ops = 0;
    int P = 32;                   // points per tile of the kernel that runs
    float* slab = nullptr;        // [slab_rows][PackLayout::total] per-CTA gradient rows
    int slab_rows = 0;
    std::vector<cudaEvent_t> ev;  // timing pairs
    size_t ev_used = 0;
    long long launches = 0;
};

struct tdq_handle {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int n_sm = 0;
    std::vector<int> layer_sizes;
    int d = 0, L = 0, HP = 0;
    long long P = 0;
    PackLayout lay{0, 0};
    float *params = nullptr, *adam_m = nullptr, *adam_v = nullptr, *mask = nullptr;
    float *wpack = nullptr, *wpack_t = nullptr;
    float* wimg = nullptr;        // tensor-core operand images of the hidden matrices (128-wide nets)
    int* flat_of_packed = nullptr;
    float *vars = nullptr, *vars_m = nullptr, *vars_v = nullptr;
    int n_vars = 0;
    float* gvec = nullptr;
    float* pinned = nullptr;      // host staging
    size_t pinned_floats = 0;
    std::vector<Term> terms;
    std::vector<Group> groups;
    bool plan_dirty = true;       // terms / point counts changed since build_plan
    float* stash = nullptr;
    size_t stash_bytes = 0;
    float* pred = nullptr;        // tdq_predict's device buffers [X d*m | u m | f m | program], kept between calls
    size_t pred_cap = 0;          // points they hold
    float* stage = nullptr;       // AoS staging for point uploads
    size_t stage_bytes = 0;
    // tdq_prefetch_term_points: uploads on a stream of their own, concurrent with the steps on `stream`
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copy = nullptr, ev_compute = nullptr;
    float* stage2 = nullptr;
    size_t stage2_bytes = 0;
    long long t_w = 0, t_lam = 0, t_vars = 0;
    // L-BFGS
    LbfgsState* lb_state = nullptr;
    float *lb_gold = nullptr, *lb_d = nullptr, *lb_S = nullptr, *lb_Y = nullptr, *lb_q = nullptr;
    float *lb_part = nullptr, *lb_hist = nullptr;     // grid-reduction partials + barrier counter; f history on the device
    size_t lb_hist_cap = 0;
    // graph-mode Newton phase (tdq_lbfgs_hz)
    float *hz_x0 = nullptr, *hz_g0 = nullptr, *hz_d = nullptr, *hz_q = nullptr, *hz_S = nullptr, *hz_Y = nullptr;
    HzScalars* hz_sc = nullptr;
    // comm
    void* comm = nullptr;
    int rank = 0, world = 1;
    // all-reduce over peer memory fused with the reduction / Adam kernel (reduce_exchange_adam_kernel)
    bool peer_ok = false;
    PeerTable peers{};
    float* peer_xbuf = nullptr;          // own exchange buffer (IPC-exported)
    unsigned* peer_flags = nullptr;      // own flag array (IPC-exported)
    void* peer_opened[2 * kMaxPeers] = {nullptr};
    int n_peer_opened = 0;
    unsigned peer_step = 0;
    int* peer_status = nullptr;          // device word raised when a wait timed out
    std::string comm_mode = "single";
    // bookkeeping
    cudaEvent_t mark0 = nullptr, mark1 = nullptr;
    bool timing = false;
    int dbg = 0;                  // TDQ_DBG developer switches (see KParams::dbg)
    long long launches = 0;
    size_t scratch_bytes = 0;
    std::string err;
};

static int round_hp(int w) { return w <= 32 ? 32 : (w <= 64 ? 64 : 128); }

static AdamDevCfg adam_dev(const tdq_adam_cfg& c, long long t) {
    // alpha = lr*sqrt(1-beta2^t)/(1-beta1^t) in fp32, as ResourceApplyAdam receives beta powers
    float b1p = powf(c.beta1, (float)t), b2p = powf(c.beta2, (float)t);
    AdamDevCfg a;
    a.alpha = c.lr * sqrtf(1.f - b2p) / (1.f - b1p);
    a.beta1 = c.beta1;
    a.beta2 = c.beta2;
    a.eps = c.eps;
    a.sign = (float)(c.sign >= 0 ? 1 : -1);
    return a;
}

static int ensure_pinned(tdq_handle* h, size_t floats) {
    if (floats <= h->pinned_floats) return 0;
    if (h->pinned) cudaFreeHost(h->pinned);
    h->pinned = nullptr;
    h->pinned_floats = 0;
    CK(h, cudaMallocHost((void**)&h->pinned, floats * sizeof(float)));
    h->pinned_floats = floats;
    return 0;
}

__global__ void aos_to_soa_kernel(const float* __restrict__ in, float* __restrict__ out, long long n, int d) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    for (int k = 0; k < d; ++k) out[(size_t)k * n + i] = in[(size_t)i * d + k];
}

static int pack_weights(tdq_handle* h, int do_adam, AdamDevCfg cfg) {
    int pp = h->lay.total();
    adam_pack_kernel<<<(pp + 255) / 256, 256, 0, h->stream>>>(pp, h->lay.off_wh(), h->HP, h->flat_of_packed, h->gvec,
                                                               h->params, h->adam_m, h->adam_v, h->mask, cfg, do_adam,
                                                               h->wpack, h->wpack_t, h->wimg);
    CK(h, cudaGetLastError());
    h->launches++;
    return 0;
}

extern "C" const char* tdq_last_error(const tdq_handle* h) { return h ? h->err.c_str() : g_last_error.c_str(); }

extern "C" int tdq_create(const int32_t* layer_sizes, int32_t n_layers, int32_t device, void* stream, tdq_handle** out) {
    tdq_handle* h0 = nullptr;
    if (!out || !layer_sizes) TDQ_FAIL(h0, "tdq_create: null argument");
    *out = nullptr;
    if (n_layers < 3) TDQ_FAIL(h0, "tdq_create: need at least [in, hidden, out]");
    if (layer_sizes[n_layers - 1] != 1) TDQ_FAIL(h0, "tdq_create: only scalar-output networks are supported (got %d outputs)", layer_sizes[n_layers - 1]);
    if (layer_sizes[0] < 1 || layer_sizes[0] > TDQ_MAX_DIMS) TDQ_FAIL(h0, "tdq_create: input dimension %d not in 1..%d", layer_sizes[0], TDQ_MAX_DIMS);
    int L = n_layers - 2;
    if (L > kMaxHidden) TDQ_FAIL(h0, "tdq_create: more than %d tanh layers", kMaxHidden);
    int wmax = 0;
    for (int i = 1; i <= L; ++i) {
        if (layer_sizes[i] < 1) TDQ_FAIL(h0, "tdq_create: bad layer width");
        wmax = layer_sizes[i] > wmax ? layer_sizes[i] : wmax;
    }
    if (wmax > 128) TDQ_FAIL(h0, "tdq_create: hidden width %d > 128 is not supported yet", wmax);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) TDQ_FAIL(h0, "tdq_create: no CUDA device (%s); this library has no CPU path", cudaGetErrorString(e));
    if (device < 0 || device >= ndev) TDQ_FAIL(h0, "tdq_create: device %d out of range (%d visible)", device, ndev);
    tdq_handle* h = new tdq_handle();
    h->device = device;
    if (cudaSetDevice(device) != cudaSuccess) { delete h; TDQ_FAIL(h0, "cudaSetDevice failed"); }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    h->n_sm = prop.multiProcessorCount;
    if (prop.major < 10) { delete h; TDQ_FAIL(h0, "tdq_create: device is sm_%d%d; this library is built for sm_100a only", prop.major, prop.minor); }
    if (stream) h->stream = (cudaStream_t)stream;
    else { cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking); h->own_stream = true; }
    { const char* e = getenv("TDQ_DBG"); h->dbg = e ? atoi(e) : 0; }
    h->layer_sizes.assign(layer_sizes, layer_sizes + n_layers);
    h->d = layer_sizes[0];
    h->L = L;
    h->HP = round_hp(wmax);
    h->lay = PackLayout{h->HP, L};
    long long P = 0;
    for (int i = 1; i < n_layers; ++i) P += (long long)layer_sizes[i - 1] * layer_sizes[i] + layer_sizes[i];
    h->P = P;
    const int pp = h->lay.total();
    // packed index -> flat index (utils.py:7-29 layout)
    std::vector<int> map(pp, -1);
    std::vector<long long> off_w(n_layers, 0), off_b(n_layers, 0);
    long long off = 0;
    for (int i = 1; i < n_layers; ++i) {
        off_w[i] = off; off += (long long)layer_sizes[i - 1] * layer_sizes[i];
        off_b[i] = off; off += layer_sizes[i];
    }
    const int HP = h->HP;
    for (int dim = 0; dim < h->d; ++dim)
        for (int nn = 0; nn < layer_sizes[1]; ++nn) map[h->lay.off_w0() + dim * HP + nn] = (int)(off_w[1] + dim * layer_sizes[1] + nn);
    for (int l = 1; l <= L; ++l)
        for (int nn = 0; nn < layer_sizes[l]; ++nn) map[h->lay.off_b() + (l - 1) * HP + nn] = (int)(off_b[l] + nn);
    for (int nn = 0; nn < layer_sizes[L]; ++nn) map[h->lay.off_wout() + nn] = (int)(off_w[L + 1] + nn);
    map[h->lay.off_sc()] = (int)off_b[L + 1];
    for (int l = 2; l <= L; ++l)
        for (int k = 0; k < layer_sizes[l - 1]; ++k)
            for (int j = 0; j < layer_sizes[l]; ++j)
                map[h->lay.off_wh() + (l - 2) * HP * HP + k * HP + j] = (int)(off_w[l] + (long long)k * layer_sizes[l] + j);
    auto dalloc = [&](void** p, size_t bytes) -> bool {
        if (cudaMalloc(p, bytes) != cudaSuccess) return false;
        cudaMemsetAsync(*p, 0, bytes, h->stream);
        return true;
    };
    bool ok = true;
    ok &= dalloc((void**)&h->params, P * 4);
    ok &= dalloc((void**)&h->adam_m, P * 4);
    ok &= dalloc((void**)&h->adam_v, P * 4);
    ok &= dalloc((void**)&h->wpack, (size_t)pp * 4);
    ok &= dalloc((void**)&h->wpack_t, (size_t)(L > 1 ? (L - 1) : 1) * HP * HP * 4);
    if (HP == 128 && L >= 2) ok &= dalloc((void**)&h->wimg, (size_t)(L - 1) * tdq::kTcImgFloats * 4);
    ok &= dalloc((void**)&h->flat_of_packed, (size_t)pp * 4);
    ok &= dalloc((void**)&h->vars, TDQ_MAX_VARS * 4);
    ok &= dalloc((void**)&h->vars_m, TDQ_MAX_VARS * 4);
    ok &= dalloc((void**)&h->vars_v, TDQ_MAX_VARS * 4);
    ok &= dalloc((void**)&h->gvec, (P + kGvecTail) * 4);
    ok &= dalloc((void**)&h->lb_state, sizeof(LbfgsState));
    if (!ok) { tdq_destroy(h); TDQ_FAIL(h0, "tdq_create: device allocation failed"); }
    cudaMemcpyAsync(h->flat_of_packed, map.data(), (size_t)pp * 4, cudaMemcpyHostToDevice, h->stream);
    cudaStreamSynchronize(h->stream);
    cudaEventCreate(&h->mark0);
    cudaEventCreate(&h->mark1);
    AdamDevCfg z{};
    if (pack_weights(h, 0, z)) { std::string m = h->err; tdq_destroy(h); TDQ_FAIL(h0, "%s", m.c_str()); }
    *out = h;
    return 0;
}

static void free_term(Term& t) {
    cudaFree(t.X); cudaFree(t.X_next); cudaFree(t.target); cudaFree(t.lam); cudaFree(t.lam_m); cudaFree(t.lam_v);
    cudaFree(t.lam_grad); cudaFree(t.f_out);
    t = Term();
}
static void free_group(Group& g) {
    cudaFree(g.ops_dev); cudaFree(g.tcops_dev); cudaFree(g.slab);
    for (auto e : g.ev) cudaEventDestroy(e);
    g = Group();
}

extern "C" int tdq_destroy(tdq_handle* h) {
    if (!h) return 0;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->copy_stream) { cudaStreamSynchronize(h->copy_stream); cudaStreamDestroy(h->copy_stream); }
    if (h->ev_copy) cudaEventDestroy(h->ev_copy);
    if (h->ev_compute) cudaEventDestroy(h->ev_compute);
    cudaFree(h->stage2);
    for (auto& t : h->terms) free_term(t);
    for (auto& g : h->groups) free_group(g);
    cudaFree(h->params); cudaFree(h->adam_m); cudaFree(h->adam_v); cudaFree(h->mask); cudaFree(h->wpack);
    cudaFree(h->wimg); cudaFree(h->wpack_t); cudaFree(h->flat_of_packed); cudaFree(h->vars); cudaFree(h->vars_m); cudaFree(h->vars_v);
    cudaFree(h->gvec); cudaFree(h->stash); cudaFree(h->stage); cudaFree(h->pred); cudaFree(h->lb_state); cudaFree(h->lb_gold); cudaFree(h->lb_d);
    cudaFree(h->lb_S); cudaFree(h->lb_Y); cudaFree(h->lb_q); cudaFree(h->lb_part); cudaFree(h->lb_hist);
    cudaFree(h->hz_x0); cudaFree(h->hz_g0); cudaFree(h->hz_d); cudaFree(h->hz_q); cudaFree(h->hz_S); cudaFree(h->hz_Y); cudaFree(h->hz_sc);
    if (h->pinned) cudaFreeHost(h->pinned);
    for (int i = 0; i < h->n_peer_opened; ++i) cudaIpcCloseMemHandle(h->peer_opened[i]);
    cudaFree(h->peer_xbuf); cudaFree(h->peer_flags); cudaFree(h->peer_status);
    if (h->comm && g_nccl.destroy) g_nccl.destroy(h->comm);
    if (h->mark0) cudaEventDestroy(h->mark0);
    if (h->mark1) cudaEventDestroy(h->mark1);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return 0;
}

extern "C" int tdq_info(const tdq_handle* h, int32_t* n_sm, int64_t* n_launches, int64_t* scratch_bytes) {
    if (!h) return 1;
    if (n_sm) *n_sm = h->n_sm;
    if (n_launches) *n_launches = h->launches;
    if (scratch_bytes) *scratch_bytes = (int64_t)h->scratch_bytes;
    return 0;
}

extern "C" int tdq_set_params(tdq_handle* h, const float* flat, int64_t n) {
    if (!h || !flat) return 1;
    if (n != h->P) TDQ_FAIL(h, "tdq_set_params: expected %lld parameters, got %lld", h->P, (long long)n);
    cudaSetDevice(h->device);
    CK(h, cudaMemcpyAsync(h->params, flat, n * 4, cudaMemcpyHostToDevice, h->stream));
    AdamDevCfg z{};
    if (pack_weights(h, 0, z)) return 1;
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int tdq_get_params(tdq_handle* h, float* flat, int64_t n) {
    if (!h || !flat) return 1;
    if (n != h->P) TDQ_FAIL(h, "tdq_get_params: expected %lld parameters, got %lld", h->P, (long long)n);
    cudaSetDevice(h->device);
    CK(h, cudaMemcpyAsync(flat, h->params, n * 4, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int tdq_set_param_mask(tdq_handle* h, const float* mask, int64_t n) {
    if (!h) return 1;
    cudaSetDevice(h->device);
    if (!mask) { cudaFree(h->mask); h->mask = nullptr; return 0; }
    if (n != h->P) TDQ_FAIL(h, "tdq_set_param_mask: expected %lld entries", h->P);
    if (!h->mask) CK(h, cudaMalloc((void**)&h->mask, n * 4));
    CK(h, cudaMemcpyAsync(h->mask, mask, n * 4, cudaMemcpyHostToDevice, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int tdq_set_vars(tdq_handle* h, const float* v, int32_t n) {
    if (!h || n < 0 || n > TDQ_MAX_VARS) TDQ_FAIL(h, "tdq_set_vars: at most %d coefficients", TDQ_MAX_VARS);
    cudaSetDevice(h->device);
    h->n_vars = n;
    if (n) CK(h, cudaMemcpyAsync(h->vars, v, n * 4, cudaMemcpyHostToDevice, h->stream));
    CK(h, cudaMemsetAsync(h->vars_m, 0, TDQ_MAX_VARS * 4, h->stream));
    CK(h, cudaMemsetAsync(h->vars_v, 0, TDQ_MAX_VARS * 4, h->stream));
    h->t_vars = 0;
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int tdq_get_vars(tdq_handle* h, float* v, int32_t n) {
    if (!h || n < 0 || n > TDQ_MAX_VARS) return 1;
    cudaSetDevice(h->device);
    if (n) CK(h, cudaMemcpyAsync(v, h->vars, n * 4, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

// validate a spec; *vout = smallest compiled configuration that covers it
static int check_spec(tdq_handle* h, const tdq_term_spec* spec, int bwd, const Variant** vout) {
    if (!spec || !spec->ops) TDQ_FAIL(h, "term spec: null");
    if (spec->n_dims != h->d) TDQ_FAIL(h, "term spec: n_dims %d does not match the network input dimension %d", spec->n_dims, h->d);
    if (spec->n_ops < 1 || spec->n_ops > TDQ_MAX_OPS) TDQ_FAIL(h, "term spec: program length %d not in 1..%d", spec->n_ops, TDQ_MAX_OPS);
    if (spec->out_slot < 0 || spec->out_slot >= spec->n_ops) TDQ_FAIL(h, "term spec: out_slot out of range");
    if (spec->f_slot >= spec->n_ops) TDQ_FAIL(h, "term spec: f_slot out of range");
    int k[3] = {0, 0, 0};
    for (int i = 0; i < spec->n_dims; ++i) {
        k[i] = spec->orders[i];
        if (k[i] < 0 || k[i] > TDQ_MAX_ORDER) TDQ_FAIL(h, "term spec: derivative order %d not in 0..%d", k[i], TDQ_MAX_ORDER);
    }
    if (spec->mixed && (spec->n_dims != 2 || k[0] < 1 || k[1] < 1 || k[0] > 2 || k[1] > 2))
        TDQ_FAIL(h, "term spec: the mixed partial u_xy is carried for two input dimensions and orders 1..2 only");
    const Variant* v = find_variant(k[0], k[1], k[2], h->HP, bwd, spec->mixed);
    if (!v) TDQ_FAIL(h, "no compiled kernel covers derivative orders (%d,%d,%d) at width %d", k[0], k[1], k[2], h->HP);
    const int n_ch = 1 + k[0] + k[1] + k[2] + (spec->mixed ? 1 : 0);
    for (int i = 0; i < spec->n_ops; ++i) {
        const tdq_op& o = spec->ops[i];
        bool unary = o.op >= TDQ_OP_NEG && o.op <= TDQ_OP_POWF;
        bool binary = o.op >= TDQ_OP_ADD && o.op <= TDQ_OP_DIV;
        if (o.op < 0 || o.op > TDQ_OP_POWF) TDQ_FAIL(h, "term spec: bad opcode %d at %d", o.op, i);
        if ((unary || binary) && (o.a < 0 || o.a >= i)) TDQ_FAIL(h, "term spec: op %d reads slot %d", i, o.a);
        if (binary && (o.b < 0 || o.b >= i)) TDQ_FAIL(h, "term spec: op %d reads slot %d", i, o.b);
        if (o.op == TDQ_OP_JET && (o.a < 0 || o.a >= n_ch)) TDQ_FAIL(h, "term spec: jet channel %d out of range", o.a);
        if (o.op == TDQ_OP_COORD && (o.a < 0 || o.a >= h->d)) TDQ_FAIL(h, "term spec: coordinate %d out of range", o.a);
        if (o.op == TDQ_OP_VAR && (o.a < 0 || o.a >= TDQ_MAX_VARS)) TDQ_FAIL(h, "term spec: var %d out of range", o.a);
    }
    *vout = v;
    return 0;
}

// a checked program whose JET channels count (dim, order) in the spec's orders `k` -> the numbering of variant v
static void remap_ops(const tdq_op* in, int n_ops, const int* k, const Variant* v, std::vector<tdq_op>* out) {
    int remap[2 + 3 * TDQ_MAX_ORDER] = {0};
    const int voff[3] = {1, 1 + v->k0, 1 + v->k0 + v->k1};
    int c = 1;
    for (int dim = 0; dim < 3; ++dim)
        for (int o = 1; o <= k[dim]; ++o) remap[c++] = voff[dim] + (o - 1);
    remap[c] = v->C - 1;                                    // the mixed partial, when the program has one: last channel
    for (int i = 0; i < n_ops; ++i) {
        tdq_op o = in[i];
        if (o.op == TDQ_OP_JET) o.a = remap[o.a];
        out->push_back(o);
    }
}

extern "C" int tdq_add_term(tdq_handle* h, const tdq_term_spec* spec, int32_t* term_id) {
    if (!h) return 1;
    cudaSetDevice(h->device);
    if ((int)h->terms.size() >= kMaxTerms) TDQ_FAIL(h, "tdq_add_term: at most %d terms", kMaxTerms);
    Term t;
    if (check_spec(h, spec, 1, &t.vb)) return 1;
    t.ops.assign(spec->ops, spec->ops + spec->n_ops);
    t.n_dims = spec->n_dims;
    for (int i = 0; i < spec->n_dims; ++i) t.orders[i] = spec->orders[i];
    t.out_slot = spec->out_slot;
    t.f_slot = spec->f_slot;
    t.pair = spec->pair;
    t.mixed = spec->mixed;
    for (auto& o : t.ops) {
        if (o.op == TDQ_OP_TARGET) t.has_target = true;
        if (o.op == TDQ_OP_LAMBDA) t.has_lam = true;
    }
    size_t smem = t.vb->smem(h->L, (int)t.ops.size());
    if (smem > 227 * 1024) TDQ_FAIL(h, "tdq_add_term: kernel needs %zu B of shared memory (> 227 KB)", smem);
    h->terms.push_back(t);
    h->plan_dirty = true;
    if (term_id) *term_id = (int)h->terms.size() - 1;
    return 0;
}

static Term* get_term(tdq_handle* h, int term) {
    if (!h || term < 0 || term >= (int)h->terms.size()) return nullptr;
    return &h->terms[term];
}

static int upload_points(tdq_handle* h, const float* X, long long n, int d, float* dst_soa) {
    // AoS host rows -> SoA device columns (models.py:49 builds one [N,1] column per variable)
    size_t bytes = (size_t)n * d * 4;
    if (bytes > h->stage_bytes) {
        cudaStreamSynchronize(h->stream);
        cudaFree(h->stage);
        h->stage = nullptr;
        h->stage_bytes = 0;
        CK(h, cudaMalloc((void**)&h->stage, bytes));
        h->stage_bytes = bytes;
    }
    CK(h, cudaMemcpyAsync(h->stage, X, bytes, cudaMemcpyHostToDevice, h->stream));
    aos_to_soa_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(h->stage, dst_soa, n, d);
    CK(h, cudaGetLastError());
    h->launches++;
    return 0;
}

extern "C" int tdq_set_term_points(tdq_handle* h, int32_t term, const float* X, int64_t n, int64_t n_global) {
    Term* t = get_term(h, term);
    if (!t) TDQ_FAIL(h, "tdq_set_term_points: bad term %d", term);
    if (n < 0 || (n > 0 && !X)) TDQ_FAIL(h, "tdq_set_term_points: bad arguments");
    if (t->pair && (n & 1)) TDQ_FAIL(h, "tdq_set_term_points: a paired term needs an even number of points");
    {
        long long units = t->pair ? n / 2 : n;   // a paired term averages over couples
        if (n_global < units || n_global <= 0) TDQ_FAIL(h, "tdq_set_term_points: n_global must be positive and >= the local count");
    }
    cudaSetDevice(h->device);
    if (n > t->cap) {
        cudaStreamSynchronize(h->stream);
        if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
        cudaFree(t->X); cudaFree(t->X_next); cudaFree(t->target); cudaFree(t->lam); cudaFree(t->lam_m); cudaFree(t->lam_v);
        cudaFree(t->lam_grad); cudaFree(t->f_out);
        t->X = t->X_next = t->target = t->lam = t->lam_m = t->lam_v = t->lam_grad = t->f_out = nullptr;
        t->next_n = -1;
        CK(h, cudaMalloc((void**)&t->X, (size_t)n * h->d * 4));
        CK(h, cudaMalloc((void**)&t->f_out, (size_t)n * 4));
        if (t->has_target) { CK(h, cudaMalloc((void**)&t->target, (size_t)n * 4)); CK(h, cudaMemsetAsync(t->target, 0, (size_t)n * 4, h->stream)); }
        if (t->has_lam) {
            CK(h, cudaMalloc((void**)&t->lam, (size_t)n * 4));
            CK(h, cudaMalloc((void**)&t->lam_m, (size_t)n * 4));
            CK(h, cudaMalloc((void**)&t->lam_v, (size_t)n * 4));
            CK(h, cudaMalloc((void**)&t->lam_grad, (size_t)n * 4));
            CK(h, cudaMemsetAsync(t->lam, 0, (size_t)n * 4, h->stream));
            CK(h, cudaMemsetAsync(t->lam_m, 0, (size_t)n * 4, h->stream));
            CK(h, cudaMemsetAsync(t->lam_v, 0, (size_t)n * 4, h->stream));
            CK(h, cudaMemsetAsync(t->lam_grad, 0, (size_t)n * 4, h->stream));
        }
        t->cap = n;
    }
    if (n != t->n) h->plan_dirty = true;       // the launch groups depend on the terms' sizes
    t->n = n;
    t->n_global = n_global;
    t->r_start = 0;
    t->r_count = -1;
    if (n == 0) return 0;
    if (upload_points(h, X, n, h->d, t->X)) return 1;
    CK(h, cudaStreamSynchronize(h->stream));   // the caller's buffer is free again on return
    return 0;
}

// Streaming a term's points from the host while the previous set is still being trained on (collocation points resampled
// between steps): the upload goes through a stream of its own into the term's shadow buffer; tdq_swap_term_points makes
// the shadow buffer the active one.  The reference would rebuild its X_f tensors (models.py:49-51) and retrace.
extern "C" int tdq_prefetch_term_points(tdq_handle* h, int32_t term, const float* X, int64_t n) {
    Term* t = get_term(h, term);
    if (!t) TDQ_FAIL(h, "tdq_prefetch_term_points: bad term %d", term);
    if (!X || n <= 0 || n != t->n) TDQ_FAIL(h, "tdq_prefetch_term_points: the count must equal the term's current point count (%lld); tdq_set_term_points resizes", t->n);
    cudaSetDevice(h->device);
    if (!h->copy_stream) {
        CK(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
        CK(h, cudaEventCreateWithFlags(&h->ev_copy, cudaEventDisableTiming));
        CK(h, cudaEventCreateWithFlags(&h->ev_compute, cudaEventDisableTiming));
    }
    const size_t bytes = (size_t)n * h->d * 4;
    if (!t->X_next) CK(h, cudaMalloc((void**)&t->X_next, (size_t)t->cap * h->d * 4));
    if (bytes > h->stage2_bytes) {
        cudaStreamSynchronize(h->copy_stream);
        cudaFree(h->stage2);
        h->stage2 = nullptr;
        h->stage2_bytes = 0;
        CK(h, cudaMalloc((void**)&h->stage2, bytes));
        h->stage2_bytes = bytes;
    }
    // whatever was enqueued on the compute stream so far may still read the shadow buffer's previous content
    CK(h, cudaEventRecord(h->ev_compute, h->stream));
    CK(h, cudaStreamWaitEvent(h->copy_stream, h->ev_compute, 0));
    CK(h, cudaMemcpyAsync(h->stage2, X, bytes, cudaMemcpyHostToDevice, h->copy_stream));
    aos_to_soa_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->copy_stream>>>(h->stage2, t->X_next, n, h->d);
    CK(h, cudaGetLastError());
    h->launches++;
    CK(h, cudaEventRecord(h->ev_copy, h->copy_stream));
    t->next_n = n;
    return 0;
}

extern "C" int tdq_swap_term_points(tdq_handle* h, int32_t term) {
    Term* t = get_term(h, term);
    if (!t) TDQ_FAIL(h, "tdq_swap_term_points: bad term %d", term);
    if (t->next_n < 0) TDQ_FAIL(h, "tdq_swap_term_points: nothing was prefetched for term %d", term);
    if (t->next_n != t->n) TDQ_FAIL(h, "tdq_swap_term_points: the term was resized after the prefetch");
    cudaSetDevice(h->device);
    CK(h, cudaEventSynchronize(h->ev_copy));            // the caller's host buffer is free again on return
    CK(h, cudaStreamWaitEvent(h->stream, h->ev_copy, 0));
    std::swap(t->X, t->X_next);
    t->next_n = -1;
    return 0;
}

// ---- launch planning: which terms share a launch, with which compiled configuration, and their scratch
static int build_plan(tdq_handle* h) {
    CK(h, cudaStreamSynchronize(h->stream));
    for (auto& g : h->groups) free_group(g);
    h->groups.clear();
    for (auto& t : h->terms) t.group = -1;
    std::vector<int> live;
    for (int i = 0; i < (int)h->terms.size(); ++i)
        if (h->terms[i].n > 0) live.push_back(i);
    auto weight = [&](int i) { return (double)h->terms[i].n * h->terms[i].vb->C; };
    while (!live.empty()) {
        // the heaviest remaining term leads a group; terms at most an eighth of its size (or tiny in absolute
        // terms) ride along -- evaluated with the leader's jet channels they still cost next to nothing, while
        // a second big term (e.g. a data-assimilation set) keeps its own launch and its own, cheaper, configuration
        int lead = 0;
        for (int j = 1; j < (int)live.size(); ++j)
            if (weight(live[j]) > weight(live[lead])) lead = j;
        Group g;
        g.terms.push_back(live[lead]);
        const long long n_lead = h->terms[live[lead]].n;
        live.erase(live.begin() + lead);
        for (size_t j = 0; j < live.size() && (int)g.terms.size() < kMaxSeg;) {
            const Term& t = h->terms[live[j]];
            if (t.n <= 4096 || t.n * 8 <= n_lead) { g.terms.push_back(live[j]); live.erase(live.begin() + j); }
            else ++j;
        }
        int k[3] = {0, 0, 0}, mixed = 0;
        for (int ti : g.terms) {
            for (int dmn = 0; dmn < 3; ++dmn) k[dmn] = h->terms[ti].orders[dmn] > k[dmn] ? h->terms[ti].orders[dmn] : k[dmn];
            mixed |= h->terms[ti].mixed;
        }
        g.vb = find_variant(k[0], k[1], k[2], h->HP, 1, mixed);
        if (!g.vb) TDQ_FAIL(h, "no compiled kernel covers derivative orders (%d,%d,%d) at width %d", k[0], k[1], k[2], h->HP);
        std::vector<tdq_op> ops;
        std::vector<TcOp> tcops;
        bool tc_ok = tc_supported(g.vb->C, h->HP, h->L) && h->wimg;
        for (int ti : g.terms) {
            const Term& t = h->terms[ti];
            Group::Prog pr{};
            pr.ops_off = (int)ops.size();
            pr.n_ops = (int)t.ops.size();
            pr.out_slot = t.out_slot;
            pr.f_slot = t.f_slot;
            std::vector<tdq_op> mine;
            remap_ops(t.ops.data(), pr.n_ops, t.orders, g.vb, &mine);
            ops.insert(ops.end(), mine.begin(), mine.end());
            TcProgram tp;
            if (tc_ok && tc_compile(mine, t.out_slot, t.f_slot, &tp)) {
                pr.tc_off = (int)tcops.size();
                pr.tc_n_ops = (int)tp.ops.size();
                pr.tc_out = tp.out;
                pr.tc_f = tp.f;
                tcops.insert(tcops.end(), tp.ops.begin(), tp.ops.end());
            } else {
                tc_ok = false;
            }
            g.prog.push_back(pr);
        }
        g.n_ops = (int)ops.size();
        g.tc_n_ops = (int)tcops.size();
        if (tc_ok && tc_smem_bytes(h->L, g.tc_n_ops) <= 227 * 1024) g.tc = find_tc_variant(g.vb->k0, g.vb->k1, g.vb->k2);
        if (g.vb->smem(h->L, g.n_ops) > 227 * 1024)
            TDQ_FAIL(h, "the merged launch needs %zu B of shared memory (> 227 KB)", g.vb->smem(h->L, g.n_ops));
        CK(h, cudaMalloc((void**)&g.ops_dev, ops.size() * sizeof(tdq_op)));
        CK(h, cudaMemcpyAsync(g.ops_dev, ops.data(), ops.size() * sizeof(tdq_op), cudaMemcpyHostToDevice, h->stream));
        if (g.tc) {
            CK(h, cudaMalloc((void**)&g.tcops_dev, (tcops.size() + 1) * sizeof(TcOp)));
            CK(h, cudaMemcpyAsync(g.tcops_dev, tcops.data(), tcops.size() * sizeof(TcOp), cudaMemcpyHostToDevice, h->stream));
        }
        CK(h, cudaStreamSynchronize(h->stream));        // the vectors are locals
        g.P = (g.tc && !(h->dbg & 16)) ? g.tc->P : g.vb->P;     // points per tile of the kernel that runs
        long long tiles = 0;
        for (int ti : g.terms) tiles += (h->terms[ti].n + g.P - 1) / g.P;
        g.slab_rows = (int)(tiles < h->n_sm ? tiles : h->n_sm);
        CK(h, cudaMalloc((void**)&g.slab, (size_t)g.slab_rows * h->lay.total() * 4));
        // per-CTA stash slab shared by all launches (they are serialised on the stream)
        size_t need = (size_t)g.slab_rows * h->L * (g.vb->C * 4) * kThreads * g.vb->tp * 4;
        if (need > h->stash_bytes) {
            cudaFree(h->stash);
            h->stash = nullptr;
            h->stash_bytes = 0;
            CK(h, cudaMalloc((void**)&h->stash, need));
            h->stash_bytes = need;
            // The stash is rewritten every tile and read back within the same tile: ask L2 to keep it
            // (persisting access-policy window) so its dirty lines are not written back to HBM.  Best
            // effort: failures only cost DRAM traffic, never correctness.
            // Only on a stream this library created: the window is a stream attribute and the set-aside a device-wide limit,
            // and a caller who handed us their own stream keeps their own L2 policy.
            if (h->own_stream && !getenv("TDQ_NO_L2_PERSIST")) {
                cudaDeviceProp prop;
                if (cudaGetDeviceProperties(&prop, h->device) == cudaSuccess && prop.persistingL2CacheMaxSize > 0) {
                    size_t want = need < (size_t)prop.persistingL2CacheMaxSize ? need : (size_t)prop.persistingL2CacheMaxSize;
                    cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want);
                    cudaStreamAttrValue av;
                    memset(&av, 0, sizeof(av));
                    size_t win = need < (size_t)prop.accessPolicyMaxWindowSize ? need : (size_t)prop.accessPolicyMaxWindowSize;
                    av.accessPolicyWindow.base_ptr = h->stash;
                    av.accessPolicyWindow.num_bytes = win;
                    av.accessPolicyWindow.hitRatio = win <= want ? 1.0f : (float)want / (float)win;
                    av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                    av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                    cudaStreamSetAttribute(h->stream, cudaStreamAttributeAccessPolicyWindow, &av);
                    cudaGetLastError();
                }
            }
        }
        const int gi = (int)h->groups.size();
        for (int ti : g.terms) h->terms[ti].group = gi;
        h->groups.push_back(g);
    }
    size_t sc = h->stash_bytes;
    for (auto& g : h->groups) sc += (size_t)g.slab_rows * h->lay.total() * 4;
    h->scratch_bytes = sc;
    h->plan_dirty = false;
    return 0;
}

extern "C" int tdq_set_term_range(tdq_handle* h, int32_t term, int64_t start, int64_t count, int64_t n_global) {
    Term* t = get_term(h, term);
    if (!t) TDQ_FAIL(h, "tdq_set_term_range: bad term %d", term);
    if (count < 0) { t->r_start = 0; t->r_count = -1; return 0; }
    if (start < 0 || start + count > t->n || n_global <= 0) TDQ_FAIL(h, "tdq_set_term_range: range outside the stored points");
    if (t->pair && ((start | count) & 1)) TDQ_FAIL(h, "tdq_set_term_range: a paired term needs an even range");
    t->r_start = start;
    t->r_count = count;
    t->r_global = n_global;
    return 0;
}

static int set_vec(tdq_handle* h, float* dst, const float* src, long long n) {
    CK(h, cudaMemcpyAsync(dst, src, (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}
static int get_vec(tdq_handle* h, float* dst, const float* src, long long n) {
    CK(h, cudaMemcpyAsync(dst, src, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int tdq_set_term_target(tdq_handle* h, int32_t term, const float* target, int64_t n) {
    Term* t = get_term(h, term);
    if (!t || !target) TDQ_FAIL(h, "tdq_set_term_target: bad arguments");
    if (!t->has_target) TDQ_FAIL(h, "tdq_set_term_target: the term's program reads no target");
    if (n != t->n) TDQ_FAIL(h, "tdq_set_term_target: expected %lld values (set the points first)", t->n);
    cudaSetDevice(h->device);
    return n ? set_vec(h, t->target, target, n) : 0;
}

extern "C" int tdq_set_term_lambda(tdq_handle* h, int32_t term, const float* lam, int64_t n) {
    Term* t = get_term(h, term);
    if (!t || !lam) TDQ_FAIL(h, "tdq_set_term_lambda: bad arguments");
    if (!t->has_lam) TDQ_FAIL(h, "tdq_set_term_lambda: the term's program reads no lambda");
    if (n != t->n) TDQ_FAIL(h, "tdq_set_term_lambda: expected %lld values (set the points first)", t->n);
    cudaSetDevice(h->device);
    if (!n) return 0;
    CK(h, cudaMemsetAsync(t->lam_m, 0, (size_t)n * 4, h->stream));
    CK(h, cudaMemsetAsync(t->lam_v, 0, (size_t)n * 4, h->stream));
    return set_vec(h, t->lam, lam, n);
}

extern "C" int tdq_get_term_lambda(tdq_handle* h, int32_t term, float* lam, int64_t n) {
    Term* t = get_term(h, term);
    if (!t || !lam || !t->has_lam || n != t->n) TDQ_FAIL(h, "tdq_get_term_lambda: bad arguments");
    cudaSetDevice(h->device);
    return n ? get_vec(h, lam, t->lam, n) : 0;
}

extern "C" int tdq_get_term_lambda_grad(tdq_handle* h, int32_t term, float* grad, int64_t n) {
    Term* t = get_term(h, term);
    if (!t || !grad || !t->has_lam || n != t->n) TDQ_FAIL(h, "tdq_get_term_lambda_grad: bad arguments");
    cudaSetDevice(h->device);
    return n ? get_vec(h, grad, t->lam_grad, n) : 0;
}

extern "C" int tdq_get_term_residual(tdq_handle* h, int32_t term, float* f, int64_t n) {
    Term* t = get_term(h, term);
    if (!t || !f || n != t->n) TDQ_FAIL(h, "tdq_get_term_residual: bad arguments");
    if (t->f_slot < 0) TDQ_FAIL(h, "tdq_get_term_residual: the term has no f_slot");
    cudaSetDevice(h->device);
    return n ? get_vec(h, f, t->f_out, n) : 0;
}

// ---- evaluate every term (one launch per group), reduce, (all-reduce) -> gvec = [flat grad | d vars | loss].
// w_cfg != null on a single GPU: the Keras-Adam update of the weights and the re-pack ride in the reduction kernel.
static int evaluate(tdq_handle* h, int lam_mode, const AdamDevCfg* lam_cfg, const AdamDevCfg* w_cfg = nullptr) {
    if (h->plan_dirty && build_plan(h)) return 1;
    SlabList sl;
    sl.n = 0;
    for (auto& g : h->groups) {
        KParams p;
        memset(&p, 0, sizeof(p));
        const int P = g.P;
        long long tiles = 0;
        int ns = 0;
        const Term* dense_lam[kMaxSeg];
        int n_dense = 0;
        for (size_t gi = 0; gi < g.terms.size(); ++gi) {
            Term& t = h->terms[g.terms[gi]];
            const long long start = t.r_count >= 0 ? t.r_start : 0;
            const long long cnt = t.r_count >= 0 ? t.r_count : t.n;
            if (cnt == 0) continue;
            Seg& q = p.seg[ns++];
            q.X = t.X + start;
            q.x_stride = t.n;
            q.target = t.target ? t.target + start : nullptr;
            q.lam = t.lam ? t.lam + start : nullptr;
            q.lam_m = t.lam_m ? t.lam_m + start : nullptr;
            q.lam_v = t.lam_v ? t.lam_v + start : nullptr;
            q.lam_grad = t.lam_grad ? t.lam_grad + start : nullptr;
            q.f_out = t.f_slot >= 0 ? t.f_out + start : nullptr;
            q.u_out = nullptr;
            q.n = cnt;
            q.tile0 = tiles;
            tiles += (cnt + P - 1) / P;
            q.inv_n = 1.0f / (float)(t.r_count >= 0 ? t.r_global : t.n_global);
            q.lam_mode = t.has_lam ? lam_mode : 0;
            if (t.has_lam && lam_mode == 2 && t.r_count >= 0 && t.r_count < t.n) {
                // minibatch: keras applies Adam to the WHOLE lambda variable, entries outside the batch see a zero
                // gradient but still move with their decaying moments (fit.py:104).  Write the batch's gradients
                // into a zeroed vector here; the dense update follows the kernel.
                q.lam_mode = 1;
                CK(h, cudaMemsetAsync(t.lam_grad, 0, (size_t)t.n * 4, h->stream));
                dense_lam[n_dense++] = &t;
            }
            q.pair = t.pair;
            const Group::Prog& pr = g.prog[gi];
            q.ops_off = pr.ops_off; q.n_ops = pr.n_ops; q.out_slot = pr.out_slot; q.f_slot = pr.f_slot;
            q.tc_off = pr.tc_off; q.tc_n_ops = pr.tc_n_ops; q.tc_out = pr.tc_out; q.tc_f = pr.tc_f;
        }
        if (ns == 0) continue;
        p.n_seg = ns;
        p.n_tiles = tiles;
        if (lam_cfg) { p.lam_adam.alpha = lam_cfg->alpha; p.lam_adam.beta1 = lam_cfg->beta1; p.lam_adam.beta2 = lam_cfg->beta2; p.lam_adam.eps = lam_cfg->eps; p.lam_adam.sign = lam_cfg->sign; }
        p.ops = g.ops_dev;
        p.n_ops = g.n_ops;
        p.vars = h->vars;
        p.n_vars = h->n_vars;
        p.d = h->d;
        p.L = h->L;
        p.wpack = h->wpack;
        p.wpack_t = h->wpack_t;
        p.wimg = h->wimg;
        p.tc_ops = g.tcops_dev;
        p.tc_n_ops = g.tc_n_ops;
        p.stash = h->stash;
        p.slab = g.slab;
        p.dbg = h->dbg;
        const bool use_tc = g.tc && !(h->dbg & 16);        // every hidden GEMM on the tensor cores
        size_t smem = use_tc ? tc_smem_bytes(h->L, g.tc_n_ops) : g.vb->smem(h->L, g.n_ops);
        bool timed = h->timing && g.ev_used + 2 <= 4096;
        if (timed) {
            while (g.ev.size() < g.ev_used + 2) { cudaEvent_t e; cudaEventCreate(&e); g.ev.push_back(e); }
            cudaEventRecord(g.ev[g.ev_used], h->stream);
        }
        int grid = (int)(tiles < g.slab_rows ? tiles : g.slab_rows);
        CK(h, use_tc ? g.tc->launch(p, grid, smem, h->stream) : g.vb->launch(p, grid, smem, h->stream));
        if (timed) { cudaEventRecord(g.ev[g.ev_used + 1], h->stream); g.ev_used += 2; }
        g.launches++;
        h->launches++;
        for (int i = 0; i < n_dense && lam_cfg; ++i) {
            const Term* t = dense_lam[i];
            adam_vec_kernel<<<(unsigned)((t->n + 255) / 256), 256, 0, h->stream>>>(t->n, t->lam_grad, t->lam, t->lam_m, t->lam_v, *lam_cfg);
            CK(h, cudaGetLastError());
            h->launches++;
        }
        sl.ptr[sl.n] = g.slab;
        sl.perm[sl.n] = use_tc ? 1 : 0;
        sl.rows[sl.n] = grid;
        sl.n++;
    }
    int pp = h->lay.total();
    AdamDevCfg z{};
    if (h->world > 1 && h->peer_ok) {
        // one kernel: local slab sums -> exchange over NVLink peer memory -> sum in rank order -> Adam + re-pack
        h->peer_step++;
        const int n_chunks = (pp + kReduceCols - 1) / kReduceCols;
        int grid = n_chunks < h->peers.max_blocks ? n_chunks : h->peers.max_blocks;
        if ((n_chunks + grid - 1) / grid > 4) TDQ_FAIL(h, "peer exchange: network too large for the resident grid");   // kMaxChunks
        reduce_exchange_adam_kernel<<<grid, dim3(kReduceCols, kReduceRows), 0, h->stream>>>(
            sl, h->peers, h->peer_step, pp, h->lay.off_sc(), h->flat_of_packed, h->gvec, (int)h->P, h->lay.off_wh(), h->HP,
            sl.n ? sl.perm[0] : 0, w_cfg ? 1 : 0, w_cfg ? *w_cfg : z, h->params, h->adam_m, h->adam_v, h->mask, h->wpack,
            h->wpack_t, h->wimg, h->peer_status);
        CK(h, cudaGetLastError());
        h->launches++;
        return 0;
    }
    const bool fuse_adam = w_cfg && h->world == 1;
    reduce_slabs_kernel<<<(pp + kReduceCols - 1) / kReduceCols, dim3(kReduceCols, kReduceRows), 0, h->stream>>>(
        sl, pp, h->lay.off_sc(), h->flat_of_packed, h->gvec, (int)h->P, h->lay.off_wh(), h->HP, sl.n ? sl.perm[0] : 0,
        fuse_adam ? 1 : 0, fuse_adam ? *w_cfg : z, h->params, h->adam_m, h->adam_v, h->mask, h->wpack, h->wpack_t, h->wimg);
    CK(h, cudaGetLastError());
    h->launches++;
    if (h->world > 1) {
        int r = g_nccl.allreduce(h->gvec, h->gvec, (size_t)(h->P + kGvecTail), /*ncclFloat*/ 7, /*ncclSum*/ 0, h->comm, h->stream);
        if (r != 0) TDQ_FAIL(h, "ncclAllReduce failed: %s", g_nccl.errstr ? g_nccl.errstr(r) : "?");
        if (w_cfg && pack_weights(h, 1, *w_cfg)) return 1;
    }
    return 0;
}

// did a peer exchange time out? (checked where the host synchronises anyway)
static int check_peer_status(tdq_handle* h) {
    if (!h->peer_ok) return 0;
    int st = 0;
    CK(h, cudaMemcpyAsync(&st, h->peer_status, 4, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    if (st) TDQ_FAIL(h, "multi-GPU exchange timed out waiting for a peer rank (is every rank running the same steps?)");
    return 0;
}

extern "C" int tdq_loss_grad(tdq_handle* h, float* loss, float* grad_flat, float* var_grads) {
    if (!h) return 1;
    cudaSetDevice(h->device);
    if (evaluate(h, 1, nullptr)) return 1;
    if (ensure_pinned(h, (size_t)h->P + kGvecTail)) return 1;
    CK(h, cudaMemcpyAsync(h->pinned, h->gvec, (size_t)(h->P + kGvecTail) * 4, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    if (check_peer_status(h)) return 1;
    if (loss) *loss = h->pinned[h->P + 8];
    if (grad_flat) memcpy(grad_flat, h->pinned, (size_t)h->P * 4);
    if (var_grads) memcpy(var_grads, h->pinned + h->P, (size_t)h->n_vars * 4);
    return 0;
}

extern "C" int tdq_adam_reset(tdq_handle* h) {
    if (!h) return 1;
    cudaSetDevice(h->device);
    CK(h, cudaMemsetAsync(h->adam_m, 0, h->P * 4, h->stream));
    CK(h, cudaMemsetAsync(h->adam_v, 0, h->P * 4, h->stream));
    CK(h, cudaMemsetAsync(h->vars_m, 0, TDQ_MAX_VARS * 4, h->stream));
    CK(h, cudaMemsetAsync(h->vars_v, 0, TDQ_MAX_VARS * 4, h->stream));
    for (auto& t : h->terms)
        if (t.has_lam && t.n) {
            CK(h, cudaMemsetAsync(t.lam_m, 0, (size_t)t.n * 4, h->stream));
            CK(h, cudaMemsetAsync(t.lam_v, 0, (size_t)t.n * 4, h->stream));
        }
    h->t_w = h->t_lam = h->t_vars = 0;
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

// one Adam step (fit.py:92-107), enqueued: evaluate, Adam on the weights / lambdas / vars; the step's loss -> *loss_host
static int adam_step_enqueue(tdq_handle* h, const tdq_adam_cfg* w, const tdq_adam_cfg* lam, const tdq_adam_cfg* vars, float* loss_host) {
    AdamDevCfg lc{};
    if (lam) { h->t_lam++; lc = adam_dev(*lam, h->t_lam); }
    AdamDevCfg wc{};
    if (w) { h->t_w++; wc = adam_dev(*w, h->t_w); }
    // gradients of every term, their reduction and -- fused into the reduction kernel on one GPU, after the
    // all-reduce otherwise -- the Adam update of the weights with the re-pack of the kernel's weight images
    if (evaluate(h, lam ? 2 : 0, lam ? &lc : nullptr, w ? &wc : nullptr)) return 1;
    if (loss_host) CK(h, cudaMemcpyAsync(loss_host, h->gvec + h->P + 8, 4, cudaMemcpyDeviceToHost, h->stream));
    if (vars && h->n_vars) {
        h->t_vars++;
        adam_vars_kernel<<<1, 32, 0, h->stream>>>(h->n_vars, h->gvec + h->P, h->vars, h->vars_m, h->vars_v, adam_dev(*vars, h->t_vars));
        CK(h, cudaGetLastError());
        h->launches++;
    }
    return 0;
}

extern "C" int tdq_adam_steps(tdq_handle* h, const tdq_adam_cfg* w, const tdq_adam_cfg* lam, const tdq_adam_cfg* vars,
                              int32_t n_steps, float* loss_hist) {
    if (!h || n_steps < 0) return 1;
    cudaSetDevice(h->device);
    if (loss_hist && ensure_pinned(h, (size_t)n_steps + (size_t)h->P + kGvecTail)) return 1;
    for (int s = 0; s < n_steps; ++s)
        if (adam_step_enqueue(h, w, lam, vars, loss_hist ? h->pinned + s : nullptr)) return 1;
    CK(h, cudaStreamSynchronize(h->stream));
    if (loss_hist && check_peer_status(h)) return 1;
    if (loss_hist) memcpy(loss_hist, h->pinned, (size_t)n_steps * 4);
    return 0;
}

extern "C" int tdq_adam_epochs(tdq_handle* h, const tdq_adam_cfg* w, const tdq_adam_cfg* lam, const tdq_adam_cfg* vars,
                               const int32_t* terms, int32_t n_terms, int64_t batch_sz, int32_t n_batches, int32_t n_epochs,
                               float* loss_hist) {
    if (!h || n_epochs < 0 || n_batches < 1 || batch_sz < 1 || n_terms < 0 || (n_terms && !terms)) TDQ_FAIL(h, "tdq_adam_epochs: bad arguments");
    cudaSetDevice(h->device);
    for (int i = 0; i < n_terms; ++i) {
        Term* t = get_term(h, terms[i]);
        if (!t) TDQ_FAIL(h, "tdq_adam_epochs: bad term %d", terms[i]);
        if ((long long)batch_sz * n_batches > t->n) TDQ_FAIL(h, "tdq_adam_epochs: %d batches of %lld exceed the %lld stored points of term %d", n_batches, (long long)batch_sz, t->n, terms[i]);
        if (t->pair && (batch_sz & 1)) TDQ_FAIL(h, "tdq_adam_epochs: a paired term needs an even batch size");
    }
    if (loss_hist && ensure_pinned(h, (size_t)n_epochs + (size_t)h->P + kGvecTail)) return 1;
    for (int e = 0; e < n_epochs; ++e)
        for (int b = 0; b < n_batches; ++b) {
            for (int i = 0; i < n_terms; ++i) {
                Term& t = h->terms[terms[i]];
                t.r_start = (long long)b * batch_sz;
                t.r_count = batch_sz;
                t.r_global = batch_sz;
            }
            // the epoch reports its last batch's loss (fit.py:109)
            if (adam_step_enqueue(h, w, lam, vars, (loss_hist && b == n_batches - 1) ? h->pinned + e : nullptr)) {
                for (int i = 0; i < n_terms; ++i) { Term& t = h->terms[terms[i]]; t.r_start = 0; t.r_count = -1; }
                return 1;
            }
        }
    for (int i = 0; i < n_terms; ++i) { Term& t = h->terms[terms[i]]; t.r_start = 0; t.r_count = -1; }
    CK(h, cudaStreamSynchronize(h->stream));
    if (loss_hist) memcpy(loss_hist, h->pinned, (size_t)n_epochs * 4);
    return 0;
}

extern "C" int tdq_lbfgs(tdq_handle* h, int32_t max_iter, float learning_rate, float* f_hist, int32_t* n_hist, int32_t* n_eval) {
    if (!h || max_iter < 0) return 1;
    cudaSetDevice(h->device);
    const int n = (int)h->P;
    if (!h->lb_gold) {
        CK(h, cudaMalloc((void**)&h->lb_gold, (size_t)n * 4));
        CK(h, cudaMalloc((void**)&h->lb_d, (size_t)n * 4));
        CK(h, cudaMalloc((void**)&h->lb_q, (size_t)n * 4));
        CK(h, cudaMalloc((void**)&h->lb_S, (size_t)n * 4 * kLbfgsHist));
        CK(h, cudaMalloc((void**)&h->lb_Y, (size_t)n * 4 * kLbfgsHist));
        CK(h, cudaMalloc((void**)&h->lb_part, (size_t)kLbfgsRounds * 2 * kLbfgsMaxCtas * 4 + 256));
        CK(h, cudaMemsetAsync(h->lb_part, 0, (size_t)kLbfgsRounds * 2 * kLbfgsMaxCtas * 4 + 256, h->stream));
    }
    if ((size_t)max_iter + 1 > h->lb_hist_cap) {
        cudaStreamSynchronize(h->stream);
        cudaFree(h->lb_hist);
        h->lb_hist = nullptr;
        CK(h, cudaMalloc((void**)&h->lb_hist, ((size_t)max_iter + 1) * 4));
        h->lb_hist_cap = (size_t)max_iter + 1;
    }
    unsigned* bar = reinterpret_cast<unsigned*>(h->lb_part + (size_t)kLbfgsRounds * 2 * kLbfgsMaxCtas);
    // CTAs of the iteration kernel: ~2k elements each, all resident at once (cooperative launch)
    int grid = n <= 8192 ? 1 : (n + kLbfgsThreads * 4 - 1) / (kLbfgsThreads * 4);     // small nets: one CTA, no grid barrier to pay
    if (grid > h->n_sm) grid = h->n_sm;
    if (grid > kLbfgsMaxCtas) grid = kLbfgsMaxCtas;
    if (grid < 1) grid = 1;
    if (ensure_pinned(h, sizeof(LbfgsState) / 4 + 16 + (size_t)max_iter + 1)) return 1;
    LbfgsState* hs = reinterpret_cast<LbfgsState*>(h->pinned);
    memset(hs, 0, sizeof(LbfgsState));                    // Struct(): every field 0 (optimizers.py:409-414)
    hs->max_iter = max_iter;
    hs->max_eval = max_iter * 1.25f;                       // :220
    CK(h, cudaMemcpyAsync(h->lb_state, hs, sizeof(LbfgsState), cudaMemcpyHostToDevice, h->stream));
    AdamDevCfg z{};
    float* loss_dev = h->gvec + h->P + 8;
    // f, g = opfunc(x)  :233
    if (evaluate(h, 0, nullptr)) return 1;
    lbfgs_check_kernel<<<1, kLbfgsThreads, 0, h->stream>>>(h->lb_state, n, h->gvec, loss_dev, h->lb_hist, 1, 1);
    CK(h, cudaGetLastError());
    h->launches++;
    // The loop body (:250-380) is enqueued without waiting: the stop tests run on the device (LbfgsState::done) and
    // every kernel of an iteration returns at once when one has fired.  The host reads the state every kPoll
    // iterations; at most kPoll - 1 evaluations at an unchanged x are wasted after a stop.
    const int kPoll = 16;
    bool done = false;
    for (int epoch = 1; epoch <= max_iter && !done; ++epoch) {
        float lr = learning_rate;
        float* xp = h->params;
        const float* gp = h->gvec;
        void* args[] = {&h->lb_state, (void*)&n, &lr, &xp, &gp, &h->lb_gold, &h->lb_d, &h->lb_S, &h->lb_Y, &h->lb_q, &h->lb_part, &bar};
        CK(h, cudaLaunchCooperativeKernel((const void*)lbfgs_iter_kernel, dim3(grid), dim3(kLbfgsThreads), args, 0, h->stream));
        h->launches++;
        if (pack_weights(h, 0, z)) return 1;
        const int evaluated = epoch != max_iter;             // :341-345: no evaluation after the last step
        if (evaluated && evaluate(h, 0, nullptr)) return 1;
        lbfgs_check_kernel<<<1, kLbfgsThreads, 0, h->stream>>>(h->lb_state, n, h->gvec, loss_dev, h->lb_hist, evaluated, 0);
        CK(h, cudaGetLastError());
        h->launches++;
        if (epoch % kPoll == 0 || epoch == max_iter) {
            CK(h, cudaMemcpyAsync(hs, h->lb_state, sizeof(LbfgsState), cudaMemcpyDeviceToHost, h->stream));
            CK(h, cudaStreamSynchronize(h->stream));
            done = hs->done != 0;
        }
    }
    CK(h, cudaMemcpyAsync(hs, h->lb_state, sizeof(LbfgsState), cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    const int nh = hs->n_hist;
    float* hist_host = h->pinned + sizeof(LbfgsState) / 4 + 16;
    if (nh > 0) {
        CK(h, cudaMemcpyAsync(hist_host, h->lb_hist, (size_t)nh * 4, cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaStreamSynchronize(h->stream));
        if (f_hist) memcpy(f_hist, hist_host, (size_t)nh * 4);
    }
    // a stop left the weights where the last accepted step put them; the kernels' weight images follow params
    if (pack_weights(h, 0, z)) return 1;
    CK(h, cudaStreamSynchronize(h->stream));
    if (n_hist) *n_hist = nh;
    if (n_eval) *n_eval = hs->n_eval;
    return 0;
}

// ---- graph-mode Newton phase: L-BFGS (10 pairs) + Hager-Zhang line search (fit.py:65-80, tfp.optimizer.lbfgs_minimize;
// the algorithm as restated in oracle/hz_lbfgs.py, whose header lists the sources and the parameters)
namespace {
struct HzPt { float c, f, df; };
struct HzSearch {
    tdq_handle* h;
    int n;
    float f0, df0, f_lim;
    int n_eval = 0, budget = 50;
    int err = 0;
    HzScalars* hs;           // pinned
    HzScalars* ds;           // device
    float* x0;               // position the search starts from
    bool wolfe(const HzPt& p) const {
        const float delta = 0.1f, sigma = 0.9f;
        const bool exact = p.f <= f0 + delta * p.c * df0 && p.df >= sigma * df0;
        const bool approx = p.f <= f_lim && (2.f * delta - 1.f) * df0 >= p.df && p.df >= sigma * df0;
        return exact || approx;
    }
    HzPt phi(float c) {
        HzPt p{c, NAN, NAN};
        AdamDevCfg z{};
        hz_trial_kernel<<<(n + 255) / 256, 256, 0, h->stream>>>(n, x0, h->lb_d, c, h->params);
        h->launches++;
        if (pack_weights(h, 0, z) || evaluate(h, 0, nullptr)) { err = 1; return p; }
        hz_dot_kernel<<<1, 1024, 0, h->stream>>>(n, h->gvec, h->lb_d, h->gvec + h->P + 8, ds);
        h->launches++;
        if (cudaMemcpyAsync(hs, ds, sizeof(HzScalars), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
            cudaStreamSynchronize(h->stream) != cudaSuccess) { err = 1; return p; }
        n_eval++;
        p.f = hs->f;
        p.df = hs->df;
        return p;
    }
    // U3: bisect [a, c] where f'(c) < 0 and f(c) > f_lim.  Returns true when a step was accepted (in *acc).
    bool u3(HzPt& a, HzPt& b, const HzPt& a_in, const HzPt& c, HzPt* acc) {
        HzPt ah = a_in, bh = c;
        while (budget > 0 && !err) {
            budget--;
            HzPt d = phi(0.5f * ah.c + 0.5f * bh.c);
            if (wolfe(d)) { a = ah; b = bh; *acc = d; return true; }
            if (d.df >= 0.f) { a = ah; b = d; return false; }
            if (d.f <= f_lim) ah = d; else bh = d;
        }
        a = ah; b = bh;
        return false;
    }
    // U0-U3; which = 0 interval unchanged / U3, 1 the trial became b, 2 the trial became a
    bool update(HzPt& a, HzPt& b, const HzPt& c, HzPt* acc, int* which) {
        *which = 0;
        if (!(a.c < c.c && c.c < b.c)) return false;
        if (wolfe(c)) { *acc = c; return true; }
        if (c.df >= 0.f) { b = c; *which = 1; return false; }
        if (c.f <= f_lim) { a = c; *which = 2; return false; }
        HzPt a0 = a;
        return u3(a, b, a0, c, acc);
    }
    static float secant(const HzPt& a, const HzPt& b) {
        const float den = b.df - a.df;
        if (den == 0.f) return 0.5f * (a.c + b.c);
        return (a.c * b.df - b.c * a.df) / den;
    }
    // the accepted step in *acc; false: no step found
    bool run(HzPt* acc) {
        HzPt c = phi(1.0f);
        while (!(isfinite(c.f) && isfinite(c.df)) && budget > 0 && !err) { budget--; c = phi(c.c * 0.1f); }
        if (err) return false;
        if (wolfe(c)) { *acc = c; return true; }
        const HzPt zero{0.f, f0, df0};
        HzPt last_ok = zero, a{}, b{};
        bool have = false;
        while (budget > 0 && !err) {                       // bracket B0-B3
            if (c.df >= 0.f) { a = last_ok; b = c; have = true; break; }
            if (c.f > f_lim) {
                if (u3(a, b, zero, c, acc)) return true;
                have = true;
                break;
            }
            last_ok = c;
            budget--;
            c = phi(c.c * 5.0f);
            if (wolfe(c)) { *acc = c; return true; }
        }
        if (!have || err) return false;
        while (budget > 0 && !err) {                       // L1-L3
            budget--;
            const float width = b.c - a.c;
            const HzPt a_in = a, b_in = b;
            const float cs = secant(a, b);
            if (a.c < cs && cs < b.c) {
                HzPt p = phi(cs);
                int which = 0;
                if (update(a, b, p, acc, &which)) return true;
                float c2 = NAN;
                if (which == 1) c2 = secant(b_in, b);
                else if (which == 2) c2 = secant(a_in, a);
                if (c2 == c2 && a.c < c2 && c2 < b.c) {
                    HzPt p2 = phi(c2);
                    if (update(a, b, p2, acc, &which)) return true;
                }
            }
            if (b.c - a.c > 0.66f * width) {
                HzPt m = phi(0.5f * (a.c + b.c));
                int which = 0;
                if (update(a, b, m, acc, &which)) return true;
            }
            if (b.c - a.c <= 0.f) break;
        }
        return false;
    }
};
}  // namespace

extern "C" int tdq_lbfgs_hz(tdq_handle* h, int32_t max_iter, float tolerance, float* f_hist, int32_t* n_hist, int32_t* n_eval,
                            int32_t* status) {
    if (!h || max_iter < 0) return 1;
    cudaSetDevice(h->device);
    const int n = (int)h->P;
    if (!h->hz_x0) {
        CK(h, cudaMalloc((void**)&h->hz_x0, (size_t)n * 4));
        CK(h, cudaMalloc((void**)&h->hz_g0, (size_t)n * 4));
        CK(h, cudaMalloc((void**)&h->hz_d, (size_t)n * 4));
        CK(h, cudaMalloc((void**)&h->hz_q, (size_t)n * 4));
        CK(h, cudaMalloc((void**)&h->hz_S, (size_t)n * 4 * kHzHist));
        CK(h, cudaMalloc((void**)&h->hz_Y, (size_t)n * 4 * kHzHist));
        CK(h, cudaMalloc((void**)&h->hz_sc, sizeof(HzScalars)));
    }
    if (ensure_pinned(h, 64)) return 1;
    HzScalars* hs = reinterpret_cast<HzScalars*>(h->pinned);
    float* const lb_d_saved = h->lb_d;
    h->lb_d = h->hz_d;                                       // HzSearch reads the direction through lb_d
    int rc = 0, nh = 0, evals = 0, st = 0;                   // st: 0 max_iter reached, 1 converged, 2 line search failed
    HzRho rho{};
    int k = 0, head = 0;
    float gamma = 1.f;
    do {
        if (evaluate(h, 0, nullptr)) { rc = 1; break; }
        evals = 1;
        if (cudaMemcpyAsync(h->hz_x0, h->params, (size_t)n * 4, cudaMemcpyDeviceToDevice, h->stream) != cudaSuccess ||
            cudaMemcpyAsync(h->hz_g0, h->gvec, (size_t)n * 4, cudaMemcpyDeviceToDevice, h->stream) != cudaSuccess ||
            cudaMemcpyAsync(&hs->f, h->gvec + h->P + 8, 4, cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
            cudaStreamSynchronize(h->stream) != cudaSuccess) { rc = 1; break; }
        float f = hs->f;
        if (f_hist) f_hist[nh] = f;
        nh++;
        for (int it = 0; it < max_iter; ++it) {
            hz_direction_kernel<<<1, 1024, 0, h->stream>>>(n, h->gvec, h->hz_S, h->hz_Y, k, head, rho, gamma, h->hz_q, h->hz_d, h->hz_sc);
            h->launches++;
            if (cudaMemcpyAsync(hs, h->hz_sc, sizeof(HzScalars), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
                cudaStreamSynchronize(h->stream) != cudaSuccess) { rc = 1; break; }
            if (hs->gmax <= tolerance) { st = 1; break; }
            HzSearch ls{h, n, f, hs->gd, f + 1e-6f * fabsf(f)};
            ls.hs = hs;
            ls.ds = h->hz_sc;
            ls.x0 = h->hz_x0;
            HzPt acc{};
            const bool ok = ls.run(&acc);
            evals += ls.n_eval;
            if (ls.err) { rc = 1; break; }
            if (!ok) {                                       // no acceptable step: back to the start of the search
                st = 2;
                AdamDevCfg z{};
                cudaMemcpyAsync(h->params, h->hz_x0, (size_t)n * 4, cudaMemcpyDeviceToDevice, h->stream);
                pack_weights(h, 0, z);
                break;
            }
            // the accepted point is the last one evaluated: params = x0 + c d, gvec = its gradient
            const int slot = (k == kHzHist) ? head : (head + k) % kHzHist;
            hz_accept_kernel<<<1, 1024, 0, h->stream>>>(n, acc.c, h->hz_d, h->gvec, h->hz_g0, h->params, h->hz_x0,
                                                        h->hz_S + (size_t)slot * n, h->hz_Y + (size_t)slot * n, h->hz_sc);
            h->launches++;
            if (cudaMemcpyAsync(hs, h->hz_sc, sizeof(HzScalars), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
                cudaStreamSynchronize(h->stream) != cudaSuccess) { rc = 1; break; }
            if (hs->ys > 0.f) {
                if (k == kHzHist) head = (head + 1) % kHzHist; else k++;
                rho.rho[slot] = 1.f / hs->ys;
                gamma = hs->ys / hs->yy;
            }
            f = acc.f;
            if (f_hist) f_hist[nh] = f;
            nh++;
        }
    } while (0);
    h->lb_d = lb_d_saved;
    if (rc) { if (h->err.empty()) h->err = "tdq_lbfgs_hz: CUDA failure"; g_last_error = h->err; return 1; }
    CK(h, cudaStreamSynchronize(h->stream));
    if (n_hist) *n_hist = nh;
    if (n_eval) *n_eval = evals;
    if (status) *status = st;
    return 0;
}

extern "C" int tdq_predict(tdq_handle* h, const tdq_term_spec* spec, const float* X, int64_t m, float* u, float* f) {
    if (!h || !X || m < 0) TDQ_FAIL(h, "tdq_predict: bad arguments");
    cudaSetDevice(h->device);
    if (m == 0) return 0;
    const Variant* v = nullptr;
    if (check_spec(h, spec, 0, &v)) return 1;
    std::vector<tdq_op> ops;
    remap_ops(spec->ops, spec->n_ops, spec->orders, v, &ops);
    if (spec->pair) TDQ_FAIL(h, "tdq_predict: paired specs are not evaluated point-wise");
    size_t smem = v->smem(h->L, (int)ops.size());
    if (smem > 227 * 1024) TDQ_FAIL(h, "tdq_predict: kernel needs %zu B of shared memory", smem);
    // device buffers live on the handle and only grow: repeated predictions (models.py:274-284 is called per plot /
    // per error evaluation) do not go through cudaMalloc / cudaFree
    if ((size_t)m > h->pred_cap) {
        cudaStreamSynchronize(h->stream);
        cudaFree(h->pred);
        h->pred = nullptr;
        h->pred_cap = 0;
        const size_t cap = ((size_t)m + 3) & ~(size_t)3;      // keeps the program records 16-byte aligned
        CK(h, cudaMalloc((void**)&h->pred, (cap * (h->d + 2)) * 4 + TDQ_MAX_OPS * sizeof(tdq_op) + (size_t)(kTcVmMaxOps + 1) * sizeof(TcOp)));
        h->pred_cap = cap;
    }
    float *Xd = h->pred, *ud = Xd + (size_t)h->pred_cap * h->d, *fd = ud + h->pred_cap;
    tdq_op* od = reinterpret_cast<tdq_op*>(fd + h->pred_cap);
    TcOp* tod = reinterpret_cast<TcOp*>(od + TDQ_MAX_OPS);
    // 128-wide nets: the tensor-core kernel in its forward-only mode (same gate as for training terms)
    const TcVariant* tcv = nullptr;
    TcProgram tprog;
    if (tc_supported(v->C, h->HP, h->L) && h->wimg && !(h->dbg & 16) && tc_compile(ops, spec->out_slot, spec->f_slot, &tprog) &&
        tc_smem_bytes(h->L, (int)tprog.ops.size()) <= 227 * 1024)
        tcv = find_tc_variant(v->k0, v->k1, v->k2);
    int rc = 0;
    do {
        cudaMemsetAsync(fd, 0, (size_t)m * 4, h->stream);
        cudaMemcpyAsync(od, ops.data(), ops.size() * sizeof(tdq_op), cudaMemcpyHostToDevice, h->stream);
        if (upload_points(h, X, m, h->d, Xd)) { rc = 1; break; }
        KParams p;
        memset(&p, 0, sizeof(p));
        const int P = tcv ? tcv->P : v->P;
        long long tiles = (m + P - 1) / P;
        Seg& q = p.seg[0];
        q.X = Xd;
        q.x_stride = m;
        q.n = m;
        q.inv_n = 1.f / (float)m;
        q.f_out = spec->f_slot >= 0 ? fd : nullptr;
        q.u_out = ud;
        q.n_ops = (int)ops.size();
        q.out_slot = spec->out_slot;
        q.f_slot = spec->f_slot;
        p.n_seg = 1;
        p.n_tiles = tiles;
        p.ops = od;
        p.n_ops = (int)ops.size();
        p.vars = h->vars;
        p.n_vars = h->n_vars;
        p.d = h->d;
        p.L = h->L;
        p.wpack = h->wpack;
        p.wpack_t = h->wpack_t;
        p.slab = nullptr;
        if (tcv) {
            cudaMemcpyAsync(tod, tprog.ops.data(), tprog.ops.size() * sizeof(TcOp), cudaMemcpyHostToDevice, h->stream);
            q.tc_off = 0; q.tc_n_ops = (int)tprog.ops.size(); q.tc_out = tprog.out; q.tc_f = tprog.f;
            p.tc_ops = tod;
            p.tc_n_ops = (int)tprog.ops.size();
            p.wimg = h->wimg;
            smem = tc_smem_bytes(h->L, p.tc_n_ops);
        }
        int grid = (int)(tiles < h->n_sm ? tiles : h->n_sm);
        cudaError_t e = tcv ? tcv->launch_fwd(p, grid, smem, h->stream) : v->launch(p, grid, smem, h->stream);
        h->launches++;
        if (e != cudaSuccess) { h->err = std::string("tdq_predict launch: ") + cudaGetErrorString(e); rc = 1; break; }
        if (u) cudaMemcpyAsync(u, ud, (size_t)m * 4, cudaMemcpyDeviceToHost, h->stream);
        if (f) cudaMemcpyAsync(f, fd, (size_t)m * 4, cudaMemcpyDeviceToHost, h->stream);
        e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) { h->err = std::string("tdq_predict: ") + cudaGetErrorString(e); rc = 1; }
    } while (0);
    if (rc) g_last_error = h->err;
    return rc;
}

extern "C" int tdq_comm_unique_id(void* out128) {
    const char* e = nccl_load();
    if (e) { g_last_error = e; return 1; }
    NcclId id;
    int r = g_nccl.get_uid(&id);
    if (r != 0) { g_last_error = "ncclGetUniqueId failed"; return 1; }
    memcpy(out128, &id, sizeof(id));
    return 0;
}

extern "C" int tdq_comm_init(tdq_handle* h, const void* nccl_unique_id, int32_t rank, int32_t world) {
    if (!h || !nccl_unique_id || world < 1 || rank < 0 || rank >= world) TDQ_FAIL(h, "tdq_comm_init: bad arguments");
    cudaSetDevice(h->device);
    if (world == 1) { h->rank = 0; h->world = 1; return 0; }
    const char* e = nccl_load();
    if (e) TDQ_FAIL(h, "tdq_comm_init: %s", e);
    NcclId id;
    memcpy(&id, nccl_unique_id, sizeof(id));
    int r = g_nccl.init_rank(&h->comm, world, id, rank);
    if (r != 0) TDQ_FAIL(h, "ncclCommInitRank failed: %s", g_nccl.errstr ? g_nccl.errstr(r) : "?");
    h->rank = rank;
    h->world = world;
    h->comm_mode = "nccl";
    // ---- exchange over peer memory: every rank exports its exchange buffer and flag array through CUDA IPC; the
    // handles travel by ncclAllGather.  Any failure (no peer access, IPC not permitted, more than kMaxPeers ranks, or
    // TDQ_NO_PEER=1) leaves the NCCL all-reduce path in place.
    const char* no_peer = getenv("TDQ_NO_PEER");
    const bool fits = (h->lay.total() + kReduceCols - 1) / kReduceCols <= 4 * h->n_sm * 4;   // chunks a resident grid can own
    if (world <= kMaxPeers && fits && g_nccl.allgather && !(no_peer && atoi(no_peer) != 0)) {
        const int pp_pad = (h->lay.total() + 63) & ~63;
        const int max_blocks = h->n_sm * 4;
        struct Hnd { cudaIpcMemHandle_t x, f; };
        Hnd* all_d = nullptr;
        std::vector<Hnd> all(world);
        bool ok = cudaMalloc((void**)&h->peer_xbuf, (size_t)2 * pp_pad * 4) == cudaSuccess &&
                  cudaMalloc((void**)&h->peer_flags, (size_t)kMaxPeers * max_blocks * 4) == cudaSuccess &&
                  cudaMalloc((void**)&h->peer_status, 4) == cudaSuccess && cudaMalloc((void**)&all_d, sizeof(Hnd) * world) == cudaSuccess;
        Hnd mine{};
        std::string why = ok ? "" : "allocation failed";
        auto fail = [&](const char* what, cudaError_t e) { ok = false; if (why.empty()) why = std::string(what) + ": " + cudaGetErrorString(e); };
        if (ok) {
            cudaMemsetAsync(h->peer_xbuf, 0, (size_t)2 * pp_pad * 4, h->stream);
            cudaMemsetAsync(h->peer_flags, 0, (size_t)kMaxPeers * max_blocks * 4, h->stream);
            cudaMemsetAsync(h->peer_status, 0, 4, h->stream);
            cudaError_t e1 = cudaIpcGetMemHandle(&mine.x, h->peer_xbuf);
            if (e1 != cudaSuccess) fail("cudaIpcGetMemHandle", e1);
            else if ((e1 = cudaIpcGetMemHandle(&mine.f, h->peer_flags)) != cudaSuccess) fail("cudaIpcGetMemHandle", e1);
        }
        // every rank takes part in the all-gather and in the vote below, whatever happened locally
        if (all_d) {
            if (!ok) memset(&mine, 0, sizeof(mine));
            cudaMemcpyAsync(all_d + rank, &mine, sizeof(Hnd), cudaMemcpyHostToDevice, h->stream);
            int rr = g_nccl.allgather(all_d + rank, all_d, sizeof(Hnd), /*ncclChar*/ 0, h->comm, h->stream);
            if (rr != 0) { ok = false; if (why.empty()) why = "ncclAllGather failed"; }
            cudaMemcpyAsync(all.data(), all_d, sizeof(Hnd) * world, cudaMemcpyDeviceToHost, h->stream);
            cudaError_t es = cudaStreamSynchronize(h->stream);
            if (es != cudaSuccess) fail("handle exchange", es);
        }
        if (ok) {
            h->peers.rank = rank; h->peers.world = world; h->peers.max_blocks = max_blocks; h->peers.pp_pad = pp_pad;
            for (int p = 0; p < world && ok; ++p) {
                if (p == rank) { h->peers.xbuf[p] = h->peer_xbuf; h->peers.flags[p] = h->peer_flags; continue; }
                Hnd zero{};
                if (memcmp(&all[p], &zero, sizeof(Hnd)) == 0) { ok = false; if (why.empty()) why = "a peer exported no handle"; break; }
                void *px = nullptr, *pf = nullptr;
                cudaError_t eo = cudaIpcOpenMemHandle(&px, all[p].x, cudaIpcMemLazyEnablePeerAccess);
                if (eo != cudaSuccess) { fail("cudaIpcOpenMemHandle", eo); break; }
                h->peer_opened[h->n_peer_opened++] = px;
                eo = cudaIpcOpenMemHandle(&pf, all[p].f, cudaIpcMemLazyEnablePeerAccess);
                if (eo != cudaSuccess) { fail("cudaIpcOpenMemHandle", eo); break; }
                h->peer_opened[h->n_peer_opened++] = pf;
                h->peers.xbuf[p] = (float*)px;
                h->peers.flags[p] = (unsigned*)pf;
            }
        }
        cudaGetLastError();
        // vote: the fused path is used only if it came up on every rank (a float SUM of 0/1 through the collective that exists)
        float* vote = reinterpret_cast<float*>(all_d);
        float v1 = ok ? 1.f : 0.f, vsum = 0.f;
        if (all_d) {
            cudaMemcpyAsync(vote, &v1, 4, cudaMemcpyHostToDevice, h->stream);
            g_nccl.allreduce(vote, vote, 1, 7, 0, h->comm, h->stream);
            cudaMemcpyAsync(&vsum, vote, 4, cudaMemcpyDeviceToHost, h->stream);
            cudaStreamSynchronize(h->stream);
            cudaFree(all_d);
        }
        h->peer_ok = ok && vsum == (float)world;
        if (h->peer_ok) h->comm_mode = "nvlink-peer (fused reduce + exchange + Adam kernel)";
        else h->comm_mode = "nccl (peer-memory path not available: " + (why.empty() ? std::string("another rank could not set it up") : why) + ")";
    }
    return 0;
}

extern "C" int tdq_comm_mode(tdq_handle* h, char* buf, int32_t cap) {
    if (!h || !buf || cap <= 0) return 1;
    snprintf(buf, (size_t)cap, "%s", h->comm_mode.c_str());
    return 0;
}

extern "C" int tdq_mark_begin(tdq_handle* h) {
    if (!h) return 1;
    cudaSetDevice(h->device);
    CK(h, cudaEventRecord(h->mark0, h->stream));
    return 0;
}
extern "C" int tdq_mark_end(tdq_handle* h, float* ms) {
    if (!h) return 1;
    cudaSetDevice(h->device);
    CK(h, cudaEventRecord(h->mark1, h->stream));
    CK(h, cudaEventSynchronize(h->mark1));
    if (ms) CK(h, cudaEventElapsedTime(ms, h->mark0, h->mark1));
    return 0;
}
extern "C" int tdq_sync(tdq_handle* h) {
    if (!h) return 1;
    cudaSetDevice(h->device);
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}
extern "C" int tdq_set_kernel_timing(tdq_handle* h, int32_t enable) {
    if (!h) return 1;
    h->timing = enable != 0;
    return 0;
}
static Group* group_of(tdq_handle* h, int32_t term) {
    Term* t = get_term(h, term);
    if (!t) return nullptr;
    if (h->plan_dirty && build_plan(h)) return nullptr;
    return t->group >= 0 ? &h->groups[t->group] : nullptr;
}

extern "C" int tdq_term_kernel(tdq_handle* h, int32_t term, char* buf, int32_t cap) {
    if (!get_term(h, term) || !buf || cap <= 0) TDQ_FAIL(h, "tdq_term_kernel: bad arguments");
    cudaSetDevice(h->device);
    Group* g = group_of(h, term);
    if (!g) TDQ_FAIL(h, "tdq_term_kernel: the term has no points on this rank");
    if (g->tc && !(h->dbg & 16))
        snprintf(buf, (size_t)cap, "%s<%d,%d,%d,%d>", g->tc->P == 16 ? "fused_jet_tc16_kernel" : "fused_jet_tc_kernel", g->tc->k0, g->tc->k1, g->tc->k2, kTcWorkerWarps);
    else
        snprintf(buf, (size_t)cap, "fused_jet_kernel<%d,%d,%d,%d,%d,true>", g->vb->k0, g->vb->k1, g->vb->k2, g->vb->hp, g->vb->tp);
    return 0;
}

// device time of the launches that evaluated `term` (together with the other terms of its launch group)
extern "C" int tdq_kernel_time(tdq_handle* h, int32_t term, int32_t reset, float* ms, int64_t* launches) {
    if (!get_term(h, term)) TDQ_FAIL(h, "tdq_kernel_time: bad term");
    cudaSetDevice(h->device);
    Group* g = group_of(h, term);
    CK(h, cudaStreamSynchronize(h->stream));
    float tot = 0.f;
    size_t used = g ? g->ev_used : 0;
    for (size_t i = 0; i + 1 < used; i += 2) {
        float e = 0.f;
        cudaEventElapsedTime(&e, g->ev[i], g->ev[i + 1]);
        tot += e;
    }
    if (ms) *ms = tot;
    if (launches) *launches = (int64_t)(used / 2);
    if (reset && g) g->ev_used = 0;
    return 0;
}
