// tdq_loss_grad_static: the loss / gradient evaluation of ONE term for callers that are not
// allowed to allocate device memory.
//
// Who needs this.  The handle API (tdq_create, tdq_set_term_points, ...) sizes its scratch with
// cudaMalloc.  Some hosts cannot do that: a plug-in running inside an application whose own
// allocator has already reserved the GPU's memory and hands out sub-buffers; a solver step
// embedded in a control loop that must never enter the driver's allocator (cudaMalloc and
// cudaFree take process-wide locks and may synchronise the whole device); a process working
// under a fixed memory budget that was carved up at start-up.  Such a caller owns device
// buffers for its inputs and outputs and passes them in; everything this function needs
// beyond them -- the padded weight matrices, the per-CTA gradient slabs and the per-CTA
// activation stash of the fused jet kernel -- lives in ONE fixed array compiled into the
// library's device image (kStaticScratchBytes below).  Every GPU that loads the library
// carries that array until the process exits.
//
// Consequences the caller must know:
//   * the first call on a device looks the array's device address up
//     (cudaGetSymbolAddress, with that device current) and remembers it; make that first call
//     during start-up, outside any stream capture;
//   * a network / term whose scratch does not fit is refused with an error text, never
//     truncated; a network that fits only partly runs on fewer CTAs than the GPU has SMs
//     (slower, same result up to summation order);
//   * all calls share the one array and find the previous call's contents in it.  The function
//     rewrites every word it later reads, so stale contents are harmless, but two calls whose
//     kernels overlap on the GPU (different streams, different threads) corrupt each other.
//     Calls issued on one stream are ordered and safe.
//   * it calls no cudaMalloc / cudaFree / cudaMallocHost and no library routine that allocates;
//     it only enqueues one memset and three kernels on the given stream and returns -- no host
//     buffer is read after the call returns (the residual program travels as a kernel argument).
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "../../include/tdq_b200.h"
#include "variants.h"

using namespace tdq;

namespace {

constexpr size_t kStaticScratchBytes = 32u << 20;                       // 32 MiB per GPU
constexpr size_t kStaticScratchFloats = kStaticScratchBytes / sizeof(float);

// The scratch itself.  In HOST code this name is only a handle for the runtime's symbol table,
// not an address a kernel could use; the device address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) float g_static_scratch[kStaticScratchFloats];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
float* g_scratch_addr[kMaxDevices] = {nullptr};

void set_err(char* err, int err_len, const char* msg) {
    if (err && err_len > 0) {
        strncpy(err, msg, (size_t)err_len - 1);
        err[err_len - 1] = 0;
    }
}

// once per device, with that device current
int scratch_of_current_device(float** out, char* err, int err_len) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        set_err(err, err_len, "tdq_loss_grad_static: no current CUDA device");
        return 1;
    }
    std::lock_guard<std::mutex> lk(g_setup_mutex);
    if (!g_scratch_addr[dev]) {
        float* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_static_scratch);
        if (e != cudaSuccess || !p) {
            char b[256];
            snprintf(b, sizeof(b), "tdq_loss_grad_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            set_err(err, err_len, b);
            return 1;
        }
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return 0;
}

// Network description small enough to travel as a kernel argument; gives packed <-> flat index
// (reference layout utils.py:7-29) without an index table in memory.
struct NetDesc {
    int n_layers, d, L, HP;
    int width[kMaxHidden + 2];
    int off_w[kMaxHidden + 2];   // flat offset of layer i's kernel (i = 1..n_layers-1)
    int off_b[kMaxHidden + 2];
};

__device__ __forceinline__ int flat_of_packed_index(const NetDesc& nd, const PackLayout& lay, int i) {
    const int HP = nd.HP, L = nd.L;
    if (i < lay.off_b()) {                        // W0 [3][HP]
        int dim = i / HP, n = i - dim * HP;
        return (dim < nd.d && n < nd.width[1]) ? nd.off_w[1] + dim * nd.width[1] + n : -1;
    }
    if (i < lay.off_wout()) {                     // biases [L][HP]
        int r = i - lay.off_b();
        int l = r / HP + 1, n = r - (l - 1) * HP;
        return n < nd.width[l] ? nd.off_b[l] + n : -1;
    }
    if (i < lay.off_sc()) {                       // w_out [HP]
        int n = i - lay.off_wout();
        return n < nd.width[L] ? nd.off_w[L + 1] + n : -1;
    }
    if (i < lay.off_wh()) return i == lay.off_sc() ? nd.off_b[L + 1] : -1;
    int r = i - lay.off_wh();
    int l = r / (HP * HP) + 2;
    int kj = r - (l - 2) * HP * HP;
    int k = kj / HP, j = kj - k * HP;
    return (k < nd.width[l - 1] && j < nd.width[l]) ? nd.off_w[l] + k * nd.width[l] + j : -1;
}

// The lowered residual program travels BY VALUE as a kernel argument (160 x 16 B = 2.5 KB, inside the 4 KB
// parameter space): no host pointer enters the stream, so nothing dangles when the caller's stack frame is gone
// and the host never waits on a staging copy.
struct ProgramArg {
    tdq_op ops[TDQ_MAX_OPS];
};

// flat parameters -> padded [in][out] matrices and their transposes (every word is written); program -> ops_dev
__global__ void static_pack_kernel(NetDesc nd, const float* __restrict__ params, float* __restrict__ wpack,
                                   float* __restrict__ wpack_t, const __grid_constant__ ProgramArg prog, int n_ops,
                                   tdq_op* __restrict__ ops_dev) {
    const PackLayout lay{nd.HP, nd.L};
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_ops) ops_dev[i] = prog.ops[i];
    if (i >= lay.total()) return;
    int fi = flat_of_packed_index(nd, lay, i);
    float v = fi >= 0 ? params[fi] : 0.f;
    wpack[i] = v;
    if (i >= lay.off_wh()) {
        int r = i - lay.off_wh();
        int l = r / (nd.HP * nd.HP);
        int kj = r - l * nd.HP * nd.HP;
        int k = kj / nd.HP, j = kj - k * nd.HP;
        wpack_t[(size_t)l * nd.HP * nd.HP + j * nd.HP + k] = v;
    }
}

// per-CTA slabs -> gvec = [flat gradient | d vars (8) | loss | 0...], rows summed in a fixed order
__global__ void static_reduce_kernel(NetDesc nd, const float* __restrict__ slab, int rows, float* __restrict__ gvec,
                                     int n_flat) {
    const PackLayout lay{nd.HP, nd.L};
    const int pp = lay.total();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pp) return;
    int fi = flat_of_packed_index(nd, lay, i);
    int sc = i - lay.off_sc();
    bool tail = sc >= 1 && sc < 10;
    if (fi < 0 && !tail) return;
    float s = 0.f;
    for (int r = 0; r < rows; ++r) s += slab[(size_t)r * pp + i];
    if (fi >= 0) gvec[fi] = s;
    else if (sc == 1) gvec[n_flat + 8] = s;
    else gvec[n_flat + (sc - 2)] = s;
}

size_t align64(size_t floats) { return (floats + 63) & ~(size_t)63; }   // 256-byte steps

}  // namespace

extern "C" int tdq_loss_grad_static(const int32_t* layer_sizes, int32_t n_layers, const tdq_term_spec* spec,
                                    const float* params_dev, const float* vars_dev, int32_t n_vars,
                                    const float* x_cols_dev, const float* target_dev, const float* lam_dev, int64_t n,
                                    int64_t n_global, float* gvec_dev, float* lam_grad_dev, void* stream, char* err,
                                    int32_t err_len) {
    char msg[320];
    cudaStream_t st = (cudaStream_t)stream;
    if (!layer_sizes || !spec || !spec->ops || !params_dev || !x_cols_dev || !gvec_dev) {
        set_err(err, err_len, "tdq_loss_grad_static: null argument");
        return 1;
    }
    if (n_layers < 3 || n_layers - 2 > kMaxHidden || layer_sizes[n_layers - 1] != 1 || layer_sizes[0] < 1 ||
        layer_sizes[0] > TDQ_MAX_DIMS) {
        set_err(err, err_len, "tdq_loss_grad_static: unsupported network shape (1..3 inputs, 1..12 tanh layers, scalar output)");
        return 1;
    }
    if (n < 1 || n_global < 1 || (spec->pair && (n & 1)) || n_vars < 0 || n_vars > TDQ_MAX_VARS) {
        set_err(err, err_len, "tdq_loss_grad_static: bad point count / n_global / n_vars");
        return 1;
    }
    NetDesc nd;
    memset(&nd, 0, sizeof(nd));
    nd.n_layers = n_layers;
    nd.d = layer_sizes[0];
    nd.L = n_layers - 2;
    int wmax = 0, off = 0;
    for (int i = 0; i < n_layers; ++i) {
        nd.width[i] = layer_sizes[i];
        if (i >= 1 && i <= nd.L) wmax = layer_sizes[i] > wmax ? layer_sizes[i] : wmax;
        if (layer_sizes[i] < 1) { set_err(err, err_len, "tdq_loss_grad_static: bad layer width"); return 1; }
        if (i >= 1) {
            nd.off_w[i] = off; off += layer_sizes[i - 1] * layer_sizes[i];
            nd.off_b[i] = off; off += layer_sizes[i];
        }
    }
    const int P = off;
    if (wmax > 128) { set_err(err, err_len, "tdq_loss_grad_static: hidden width above 128 is not supported"); return 1; }
    nd.HP = wmax <= 32 ? 32 : (wmax <= 64 ? 64 : 128);
    const PackLayout lay{nd.HP, nd.L};

    // ---- the term: pick the compiled kernel, renumber the program's jet channels for it (on the stack)
    if (spec->n_dims != nd.d || spec->n_ops < 1 || spec->n_ops > TDQ_MAX_OPS || spec->out_slot < 0 ||
        spec->out_slot >= spec->n_ops || spec->f_slot >= spec->n_ops) {
        set_err(err, err_len, "tdq_loss_grad_static: malformed term spec");
        return 1;
    }
    int k[3] = {0, 0, 0};
    for (int i = 0; i < nd.d; ++i) {
        k[i] = spec->orders[i];
        if (k[i] < 0 || k[i] > TDQ_MAX_ORDER) { set_err(err, err_len, "tdq_loss_grad_static: derivative order out of range"); return 1; }
    }
    if (spec->mixed && (nd.d != 2 || k[0] < 1 || k[1] < 1 || k[0] > 2 || k[1] > 2)) { set_err(err, err_len, "tdq_loss_grad_static: the mixed partial needs two input dimensions and orders 1..2"); return 1; }
    const Variant* v = find_variant(k[0], k[1], k[2], nd.HP, 1, spec->mixed);
    if (!v) { set_err(err, err_len, "tdq_loss_grad_static: no compiled kernel covers these derivative orders"); return 1; }
    const int n_ch = 1 + k[0] + k[1] + k[2] + (spec->mixed ? 1 : 0);
    int remap[2 + 3 * TDQ_MAX_ORDER];
    {
        const int voff[3] = {1, 1 + v->k0, 1 + v->k0 + v->k1};
        remap[0] = 0;
        int c = 1;
        for (int dim = 0; dim < 3; ++dim)
            for (int o = 1; o <= k[dim]; ++o) remap[c++] = voff[dim] + (o - 1);
        remap[c] = v->C - 1;       // the mixed partial, when the program reads one
    }
    ProgramArg lowered;
    bool reads_target = false, reads_lam = false, reads_var = false;
    for (int i = 0; i < spec->n_ops; ++i) {
        tdq_op o = spec->ops[i];
        const bool unary = o.op >= TDQ_OP_NEG && o.op <= TDQ_OP_POWF, binary = o.op >= TDQ_OP_ADD && o.op <= TDQ_OP_DIV;
        bool bad = o.op < 0 || o.op > TDQ_OP_POWF || ((unary || binary) && (o.a < 0 || o.a >= i)) ||
                   (binary && (o.b < 0 || o.b >= i)) || (o.op == TDQ_OP_JET && (o.a < 0 || o.a >= n_ch)) ||
                   (o.op == TDQ_OP_COORD && (o.a < 0 || o.a >= nd.d)) || (o.op == TDQ_OP_VAR && (o.a < 0 || o.a >= n_vars));
        if (bad) {
            snprintf(msg, sizeof(msg), "tdq_loss_grad_static: malformed program at op %d", i);
            set_err(err, err_len, msg);
            return 1;
        }
        if (o.op == TDQ_OP_JET) o.a = remap[o.a];
        if (o.op == TDQ_OP_TARGET) reads_target = true;
        if (o.op == TDQ_OP_LAMBDA) reads_lam = true;
        if (o.op == TDQ_OP_VAR) reads_var = true;
        lowered.ops[i] = o;
    }
    if (reads_var && !vars_dev) {
        set_err(err, err_len, "tdq_loss_grad_static: the program reads a trainable coefficient (VAR) but vars_dev is NULL");
        return 1;
    }
    if ((reads_target && !target_dev) || (reads_lam && !lam_dev)) {
        set_err(err, err_len, "tdq_loss_grad_static: the program reads a target / lambda but no device vector was passed");
        return 1;
    }
    const size_t smem = v->smem(nd.L, spec->n_ops);
    if (smem > 227 * 1024) { set_err(err, err_len, "tdq_loss_grad_static: kernel shared memory above 227 KB"); return 1; }

    // ---- carve the static array: [wpack | wpack_t | program | slabs (grid rows) | stash (grid rows)]
    int dev = 0, n_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    const size_t pp = (size_t)lay.total();
    const size_t f_wpack = align64(pp);
    const size_t f_wt = align64((size_t)(nd.L > 1 ? nd.L - 1 : 1) * nd.HP * nd.HP);
    const size_t f_ops = align64((size_t)spec->n_ops * (sizeof(tdq_op) / sizeof(float)));
    const size_t fixed = f_wpack + f_wt + f_ops;
    const size_t per_cta_slab = align64(pp);
    const size_t per_cta_stash = (size_t)nd.L * (v->C * 4) * kThreads * v->tp;
    // slabs must be contiguous rows of exactly pp floats for the reduction, so budget with pp and align the block
    const long long tiles = (n + v->P - 1) / v->P;
    long long grid = tiles < n_sm ? tiles : n_sm;
    (void)per_cta_slab;
    while (grid >= 1 && fixed + align64((size_t)grid * pp) + (size_t)grid * per_cta_stash > kStaticScratchFloats) --grid;
    if (grid < 1) {
        snprintf(msg, sizeof(msg),
                 "tdq_loss_grad_static: this network needs %zu bytes of scratch for a single CTA; the built-in array holds %zu",
                 (fixed + align64(pp) + per_cta_stash) * sizeof(float), kStaticScratchBytes);
        set_err(err, err_len, msg);
        return 1;
    }
    float* base = nullptr;
    if (scratch_of_current_device(&base, err, err_len)) return 1;
    float* wpack = base;
    float* wpack_t = wpack + f_wpack;
    tdq_op* ops_dev = reinterpret_cast<tdq_op*>(wpack_t + f_wt);
    float* slab = wpack_t + f_wt + f_ops;
    float* stash = slab + align64((size_t)grid * pp);

    cudaError_t e;
#define SCK(call)                                                                                        \
    do {                                                                                                 \
        e = (call);                                                                                      \
        if (e != cudaSuccess) {                                                                          \
            snprintf(msg, sizeof(msg), "tdq_loss_grad_static: %s failed: %s", #call, cudaGetErrorString(e)); \
            set_err(err, err_len, msg);                                                                  \
            return 1;                                                                                    \
        }                                                                                                \
    } while (0)

    // weights and program into the scratch (the program rides in the kernel's parameter space, see ProgramArg)
    static_pack_kernel<<<(unsigned)(((pp > (size_t)spec->n_ops ? pp : (size_t)spec->n_ops) + 255) / 256), 256, 0, st>>>(nd, params_dev, wpack, wpack_t, lowered, spec->n_ops, ops_dev);
    SCK(cudaGetLastError());

    KParams p;
    memset(&p, 0, sizeof(p));
    Seg& q = p.seg[0];
    q.X = x_cols_dev;
    q.x_stride = n;
    q.target = target_dev;
    q.lam = const_cast<float*>(lam_dev);       // read only: lam_mode <= 1 never writes lambda
    q.lam_grad = lam_grad_dev;
    q.n = n;
    q.inv_n = 1.0f / (float)n_global;
    q.lam_mode = (reads_lam && lam_grad_dev) ? 1 : 0;
    q.n_ops = spec->n_ops;
    q.out_slot = spec->out_slot;
    q.f_slot = -1;
    q.pair = spec->pair;
    p.n_seg = 1;
    p.n_tiles = tiles;
    p.ops = ops_dev;
    p.n_ops = spec->n_ops;
    p.vars = vars_dev ? vars_dev : wpack;      // NULL only when the program has no VAR op (checked above): never read then
    p.n_vars = n_vars;
    p.d = nd.d;
    p.L = nd.L;
    p.wpack = wpack;
    p.wpack_t = wpack_t;
    p.stash = stash;
    p.slab = slab;
    SCK(v->launch(p, (int)grid, smem, st));

    // the output's tail (d vars, loss, padding) must not keep what the caller's buffer held before
    SCK(cudaMemsetAsync(gvec_dev + P, 0, 16 * sizeof(float), st));
    static_reduce_kernel<<<(unsigned)((pp + 255) / 256), 256, 0, st>>>(nd, slab, (int)grid, gvec_dev, P);
    SCK(cudaGetLastError());
#undef SCK
    return 0;
}
