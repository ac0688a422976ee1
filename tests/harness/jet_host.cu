// CPU harness (test infrastructure): runs the SAME jet_math.cuh / vm.cuh source the CUDA kernels
// use, serially on the host, for a small tanh-MLP -- so the jet algebra and the program
// interpreter are checked against autograd without a GPU.  Not part of the product.
//
// stdin (binary): int32 cfg[3] (K0,K1,K2), d, n_layers, layer_sizes[n_layers], n, n_ops, out_slot;
//                 float w[P]; float X[n][d]; float target[n]; float lam[n]; float vars[8]; tdq_op ops[n_ops]
// stdout (binary): float jets[n][C]; float out[n]; float grad[P]; float lam_bar[n]; float var_bar[8]
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include "../../tensordiffeq_b200/csrc/jet_math.cuh"
#include "../../tensordiffeq_b200/csrc/vm.cuh"
#include "../../tensordiffeq_b200/csrc/vm_poly.cuh"

using namespace tdq;

// TDQ_HARNESS_F2: push the jet formulas through their packed instantiation (value type F2 = two points at a time, what
// the tensor-core kernel's worker warps run); the second half carries the same point and must agree with the first.
static bool g_f2 = false;
template <class J, int C>
static void recompute_any(const float* s, float* a) {
    if (!g_f2) { recompute_any<J, C>(s, a); return; }
    F2 s2[C], a2[C];
    for (int c = 0; c < C; ++c) s2[c] = F2(s[c], s[c]);
    J::recompute(s2, a2);
    for (int c = 0; c < C; ++c) { if (a2[c].x != a2[c].y) { fprintf(stderr, "F2 halves differ\n"); exit(4); } a[c] = a2[c].x; }
}
template <class J, int C>
static void bwd_any(const float* s, const float* ab, float* zb) {
    if (!g_f2) { J::bwd(s, ab, zb); return; }
    F2 s2[C], b2[C], z2[C];
    for (int c = 0; c < C; ++c) { s2[c] = F2(s[c], s[c]); b2[c] = F2(ab[c], ab[c]); }
    J::bwd(s2, b2, z2);
    for (int c = 0; c < C; ++c) { if (z2[c].x != z2[c].y) { fprintf(stderr, "F2 halves differ\n"); exit(4); } zb[c] = z2[c].x; }
}

template <int K0, int K1, int K2>
static void run(int d, const std::vector<int>& ls, int n, const std::vector<float>& w, const std::vector<float>& X,
                const std::vector<float>& target, const std::vector<float>& lam, const float* vars,
                const std::vector<tdq_op>& ops, int out_slot) {
    using J = JetCfg<K0, K1, K2>;
    constexpr int C = J::C;
    const int nl = (int)ls.size();
    const int L = nl - 2;
    std::vector<size_t> offW(nl), offB(nl);
    size_t off = 0;
    for (int i = 1; i < nl; ++i) { offW[i] = off; off += (size_t)ls[i - 1] * ls[i]; offB[i] = off; off += ls[i]; }
    const size_t P = off;
    std::vector<float> jets((size_t)n * C), outv(n), grad(P, 0.f), lamb(n, 0.f);
    float varb[8] = {0};
    const int offs[3] = {J::OFF0, J::OFF1, J::OFF2};
    const int ks[3] = {K0, K1, K2};
    for (int p = 0; p < n; ++p) {
        std::vector<std::vector<float>> st(L + 1), act(L + 1);   // [layer][neuron*C + c]
        // layer 1
        {
            int H = ls[1];
            st[1].assign((size_t)H * C, 0.f); act[1].assign((size_t)H * C, 0.f);
            for (int j = 0; j < H; ++j) {
                float z[C]; for (int c = 0; c < C; ++c) z[c] = 0.f;
                z[0] = w[offB[1] + j];
                for (int i = 0; i < d; ++i) {
                    float wv = w[offW[1] + (size_t)i * H + j];
                    z[0] += X[(size_t)p * d + i] * wv;
                    if (ks[i] > 0) z[offs[i]] = wv;
                }
                float s[C], a[C];
                s[0] = g_f2 ? tanh_acc(F2(z[0], z[0])).x : tanh_acc(z[0]); for (int c = 1; c < C; ++c) s[c] = z[c];
                recompute_any<J, C>(s, a);
                for (int c = 0; c < C; ++c) { st[1][(size_t)j * C + c] = s[c]; act[1][(size_t)j * C + c] = a[c]; }
            }
        }
        for (int l = 2; l <= L; ++l) {
            int Hi = ls[l - 1], H = ls[l];
            st[l].assign((size_t)H * C, 0.f); act[l].assign((size_t)H * C, 0.f);
            for (int j = 0; j < H; ++j) {
                float z[C]; for (int c = 0; c < C; ++c) z[c] = 0.f;
                z[0] = w[offB[l] + j];
                for (int k = 0; k < Hi; ++k) {
                    float wv = w[offW[l] + (size_t)k * H + j];
                    for (int c = 0; c < C; ++c) z[c] = fmaf(act[l - 1][(size_t)k * C + c], wv, z[c]);
                }
                float s[C], a[C];
                s[0] = g_f2 ? tanh_acc(F2(z[0], z[0])).x : tanh_acc(z[0]); for (int c = 1; c < C; ++c) s[c] = z[c];
                recompute_any<J, C>(s, a);
                for (int c = 0; c < C; ++c) { st[l][(size_t)j * C + c] = s[c]; act[l][(size_t)j * C + c] = a[c]; }
            }
        }
        float u[C]; for (int c = 0; c < C; ++c) u[c] = 0.f;
        int HL = ls[L];
        for (int k = 0; k < HL; ++k) for (int c = 0; c < C; ++c) u[c] = fmaf(act[L][(size_t)k * C + c], w[offW[L + 1] + k], u[c]);
        u[0] += w[offB[L + 1]];
        for (int c = 0; c < C; ++c) jets[(size_t)p * C + c] = u[c];
        // program
        float coord[3] = {0, 0, 0};
        for (int i = 0; i < d; ++i) coord[i] = X[(size_t)p * d + i];
        VmIn in; in.jets = u; in.coord = coord; in.target = target[p]; in.lam = lam[p]; in.vars = vars;
        float ub[C]; for (int c = 0; c < C; ++c) ub[c] = 0.f;
        if (getenv("TDQ_HARNESS_POLY")) {
            // the compiled form of the program (vm_poly.h), as the tensor-core kernel runs it
            TcProgram prog;
            if (!tc_compile(ops, out_slot, -1, &prog)) { fprintf(stderr, "tc_compile failed\n"); exit(3); }
            const int n_ent = kTcVmInputs + (int)prog.ops.size();
            std::vector<float> V(n_ent, 0.f), A(n_ent, 0.f);
            for (int c = 0; c < C; ++c) V[c] = u[c];
            for (int i = 0; i < d; ++i) V[kTcVmCoord + i] = coord[i];
            V[kTcVmTarget] = target[p];
            V[kTcVmLambda] = lam[p];
            for (int v = 0; v < 8; ++v) V[kTcVmVars + v] = vars[v];
            tc_vm_forward<1>(prog.ops.data(), (int)prog.ops.size(), V.data());
            outv[p] = V[prog.out];
            A[prog.out] = 1.f;
            tc_vm_reverse<1>(prog.ops.data(), (int)prog.ops.size(), V.data(), A.data());
            for (int c = 0; c < C; ++c) ub[c] = A[c];
            lamb[p] = A[kTcVmLambda];
            for (int v = 0; v < 8; ++v) varb[v] += A[kTcVmVars + v];
        } else {
            float slot[TDQ_MAX_OPS], adj[TDQ_MAX_OPS];
            vm_forward(ops.data(), (int)ops.size(), in, slot);
            outv[p] = slot[out_slot];
            for (size_t i = 0; i < ops.size(); ++i) adj[i] = 0.f;
            adj[out_slot] = 1.f;
            lamb[p] = vm_reverse(ops.data(), (int)ops.size(), slot, adj, ub, varb);
        }
        // backward
        std::vector<float> ab((size_t)HL * C);
        grad[offB[L + 1]] += ub[0];
        for (int k = 0; k < HL; ++k) {
            float g = 0.f;
            for (int c = 0; c < C; ++c) { g += act[L][(size_t)k * C + c] * ub[c]; ab[(size_t)k * C + c] = ub[c] * w[offW[L + 1] + k]; }
            grad[offW[L + 1] + k] += g;
        }
        for (int l = L; l >= 1; --l) {
            int H = ls[l];
            std::vector<float> zb((size_t)H * C);
            for (int j = 0; j < H; ++j) {
                bwd_any<J, C>(&st[l][(size_t)j * C], &ab[(size_t)j * C], &zb[(size_t)j * C]);
                grad[offB[l] + j] += zb[(size_t)j * C];
            }
            if (l == 1) {
                for (int i = 0; i < d; ++i)
                    for (int j = 0; j < H; ++j) {
                        float g = X[(size_t)p * d + i] * zb[(size_t)j * C];
                        if (ks[i] > 0) g += zb[(size_t)j * C + offs[i]];
                        grad[offW[1] + (size_t)i * H + j] += g;
                    }
                break;
            }
            int Hi = ls[l - 1];
            std::vector<float> abp((size_t)Hi * C, 0.f);
            for (int k = 0; k < Hi; ++k)
                for (int j = 0; j < H; ++j) {
                    float wv = w[offW[l] + (size_t)k * H + j];
                    float g = 0.f;
                    for (int c = 0; c < C; ++c) {
                        g += act[l - 1][(size_t)k * C + c] * zb[(size_t)j * C + c];
                        abp[(size_t)k * C + c] += zb[(size_t)j * C + c] * wv;
                    }
                    grad[offW[l] + (size_t)k * H + j] += g;
                }
            ab.swap(abp);
        }
    }
    fwrite(jets.data(), 4, jets.size(), stdout);
    fwrite(outv.data(), 4, outv.size(), stdout);
    fwrite(grad.data(), 4, grad.size(), stdout);
    fwrite(lamb.data(), 4, lamb.size(), stdout);
    fwrite(varb, 4, 8, stdout);
}

int main() {
    g_f2 = getenv("TDQ_HARNESS_F2") != nullptr;
    int hdr[5];
    if (fread(hdr, 4, 5, stdin) != 5) return 2;
    int K0 = hdr[0], K1 = hdr[1], K2 = hdr[2], d = hdr[3], nl = hdr[4];
    std::vector<int> ls(nl);
    if (fread(ls.data(), 4, nl, stdin) != (size_t)nl) return 2;
    int tail[3];
    if (fread(tail, 4, 3, stdin) != 3) return 2;
    int n = tail[0], n_ops = tail[1], out_slot = tail[2];
    size_t P = 0;
    for (int i = 1; i < nl; ++i) P += (size_t)ls[i - 1] * ls[i] + ls[i];
    std::vector<float> w(P), X((size_t)n * d), target(n), lam(n);
    float vars[8];
    std::vector<tdq_op> ops(n_ops);
    if (fread(w.data(), 4, P, stdin) != P) return 2;
    if (fread(X.data(), 4, X.size(), stdin) != X.size()) return 2;
    if (fread(target.data(), 4, n, stdin) != (size_t)n) return 2;
    if (fread(lam.data(), 4, n, stdin) != (size_t)n) return 2;
    if (fread(vars, 4, 8, stdin) != 8) return 2;
    if (fread(ops.data(), sizeof(tdq_op), n_ops, stdin) != (size_t)n_ops) return 2;
    int code = K0 * 100 + K1 * 10 + K2;
    switch (code) {
        case 0: run<0, 0, 0>(d, ls, n, w, X, target, lam, vars, ops, out_slot); break;
        case 110: run<1, 1, 0>(d, ls, n, w, X, target, lam, vars, ops, out_slot); break;
        case 210: run<2, 1, 0>(d, ls, n, w, X, target, lam, vars, ops, out_slot); break;
        case 220: run<2, 2, 0>(d, ls, n, w, X, target, lam, vars, ops, out_slot); break;
        case 410: run<4, 1, 0>(d, ls, n, w, X, target, lam, vars, ops, out_slot); break;
        case 222: run<2, 2, 2>(d, ls, n, w, X, target, lam, vars, ops, out_slot); break;
        case 310: run<3, 1, 0>(d, ls, n, w, X, target, lam, vars, ops, out_slot); break;
        default: return 3;
    }
    return 0;
}
