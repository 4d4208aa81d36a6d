// C-ABI entry points that orchestrate whole layers: the depth loops of the graph LSTM / GRU
// (ggpm/rnn.py:41-50, 96-108) and their back-propagation through time, plus library-wide state.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <initializer_list>

#include "common.cuh"
#include "kernels.h"

namespace ggpm {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};
static std::atomic<int> g_engine{GGPM_ENGINE_AUTO};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_launch(const char* what, int n_kernels) {
    g_launches.fetch_add(n_kernels, std::memory_order_relaxed);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return 3;
    }
    return 0;
}

int sm_count() {
    static int cached = 0;
    if (cached > 0) return cached;
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
        n = 148;
    cached = n;
    return n;
}

int gemm_engine() { return g_engine.load(std::memory_order_relaxed); }

bool pdl_enabled() {
    static int on = -1;
    if (on < 0) {
        const char* e = getenv("GGPM_PDL");
        on = (e != nullptr && e[0] == '0') ? 0 : 1;
    }
    return on != 0;
}

// ---- side stream of the backward depth loop ------------------------------------------------------------
// The weight-gradient GEMM of a backward round has no consumer until the loop ends, while everything else of the
// round sits on the critical path (state-gradient GEMM -> cell kernel of the next round).  It is therefore issued on
// a second stream, forked and joined with events, so that this tensor-bound GEMM runs under the HBM-bound cell
// kernel of the next round.  One stream + two events per device, created on first use.  Opt-in (GGPM_WGRAD_OVERLAP=1):
// the GEMM CTAs own most of an SM's registers, so the measured gain decides the default.
struct SideStream { cudaStream_t s; cudaEvent_t fork, join; bool ok; };
static SideStream* side_stream() {
    static SideStream side[64];
    static std::atomic<int> state[64];          // 0 = untouched, 1 = being created, 2 = ready
    static int enabled = -1;
    if (enabled < 0) {
        const char* e = getenv("GGPM_WGRAD_OVERLAP");
        enabled = (e != nullptr && e[0] == '1') ? 1 : 0;
    }
    int dev = 0;
    if (!enabled || cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    int expect = 0;
    if (state[dev].compare_exchange_strong(expect, 1)) {
        SideStream& q = side[dev];
        q.ok = cudaStreamCreateWithFlags(&q.s, cudaStreamNonBlocking) == cudaSuccess &&
               cudaEventCreateWithFlags(&q.fork, cudaEventDisableTiming) == cudaSuccess &&
               cudaEventCreateWithFlags(&q.join, cudaEventDisableTiming) == cudaSuccess;
        state[dev].store(2);
    }
    while (state[dev].load() != 2) {}
    return side[dev].ok ? &side[dev] : nullptr;
}

// ---- per-launch event timing -------------------------------------------------------------------------
constexpr int kProfMax = 1 << 16;
static bool g_prof_on = false;
static int g_prof_n = 0;
static cudaEvent_t* g_prof_ev = nullptr;        // 2 * kProfMax events, created on first enable
static int g_prof_cls[kProfMax];
static double g_prof_work[kProfMax];

ProfScope::ProfScope(int cls, cudaStream_t stream, double work) : slot(-1), st(stream) {
    if (!g_prof_on || g_prof_n >= kProfMax) return;
    slot = g_prof_n++;
    g_prof_cls[slot] = cls;
    g_prof_work[slot] = work;
    cudaEventRecord(g_prof_ev[2 * slot], st);
}
ProfScope::~ProfScope() {
    if (slot >= 0) cudaEventRecord(g_prof_ev[2 * slot + 1], st);
}

// C = A[M,K] * W[N,K]^T with both operands K-contiguous
static int gemm_nt(const float* A, int lda, const float* W, int ldw, float* C, int ldc, int M, int N, int K, float beta,
                   const float* bias, int act, int zero_row0, cudaStream_t st) {
    GemmArgs g;
    g.A = A; g.sam = lda; g.sak = 1;
    g.B = W; g.sbk = 1; g.sbn = ldw;
    g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K; g.beta = beta; g.bias = bias; g.act = act; g.zero_row0 = zero_row0;
    return launch_gemm_nt(g, st);
}
// C = A[M,K] * B[K,N] with B stored row-major [K, N]  (input gradients: dX = dY * W)
static int gemm_nn(const float* A, int lda, const float* B, int ldb, float* C, int ldc, int M, int N, int K, float beta,
                   cudaStream_t st) {
    GemmArgs g;
    g.A = A; g.sam = lda; g.sak = 1;
    g.B = B; g.sbk = ldb; g.sbn = 1;
    g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K; g.beta = beta;
    return launch_gemm_nt(g, st);
}
// C[M,N] += A^T * B with A stored [K, M], B stored [K, N]  (weight gradients, K = number of messages)
static int gemm_tn_acc(const float* A, int lda, const float* B, int ldb, float* C, int ldc, int M, int N, int K,
                       cudaStream_t st) {
    GemmArgs g;
    g.A = A; g.sam = 1; g.sak = lda;
    g.B = B; g.sbk = ldb; g.sbn = 1;
    g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K; g.beta = 1.f;
    // enough K-slices to fill the machine: tiles(M,N) * splits >= ~2 waves
    int tiles = ceil_div(M, 128) * ceil_div(N, 128);
    int splits = ceil_div(2 * sm_count(), tiles);
    int max_splits = ceil_div(K, 256);
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    g.split_k = splits;
    return launch_gemm_nt(g, st);
}

// enough K-slices to fill the machine with 128 x 256 tiles, each slice at least 256 deep
static int wgrad_splits(int M, int N, int K) {
    int tiles = ceil_div(M, 128) * ceil_div(N, 256);
    int splits = ceil_div(sm_count(), tiles);
    int max_splits = ceil_div(K, 256);
    if (splits > max_splits) splits = max_splits;
    return splits < 1 ? 1 : splits;
}

// the fused tcgen05 layer path needs the tensor-core engine and 16-byte aligned tables
static bool tc_layer_ok(int Hp, std::initializer_list<const void*> ptrs) {
    if (gemm_engine() == GGPM_ENGINE_SIMT_FP32 || Hp % 32 != 0) return false;
    for (const void* q : ptrs)
        if (((uintptr_t)q & 15) != 0) return false;
    return true;
}

// f16x3 engine workspace inside the caller's w_split buffer (8 Hp^2 floats): fp16 hi plane, fp16 lo plane
// (4 Hp^2 halves each), then scalars: [0] max|W|, [1] weight scale, [2] max neighbour count (int),
// [16 + 16 t + slot] max |gate-gradient term| of backward round t
struct F16Work { void* hi; void* lo; float* scal; };
static F16Work f16_work(float* w_split, int Hp) {
    F16Work w;
    const size_t plane_floats = (size_t)2 * Hp * Hp;      // 4 Hp^2 halves
    w.hi = w_split;
    w.lo = w_split ? w_split + plane_floats : nullptr;
    w.scal = w_split ? w_split + 2 * plane_floats : nullptr;
    return w;
}
static bool f16_scalars_fit(int Hp, int depth) { return (size_t)16 + 16 * (size_t)(depth + 1) <= (size_t)4 * Hp * Hp; }
// layers with at least this many message rows form S with a gather kernel and run the forward gate GEMM from planes
static int sgather_min_rows() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("GGPM_SGATHER_MIN_ROWS");
        v = e != nullptr ? atoi(e) : 32768;
    }
    return v;
}
// GGPM_FUSED_GATES=1: on the plane-fed path the forward gate GEMM applies the x-projection, the gate non-linearities and the
// row-0 mask in its epilogue (GGPM_ACT_LSTM_GATES) and the history keeps the activated gates; the cell kernels then only
// evaluate the per-neighbour forget gate.  Read by the forward and the backward call alike.
static bool fused_gates() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("GGPM_FUSED_GATES");
        v = (e != nullptr && e[0] == '1') ? 1 : 0;
    }
    return v != 0;
}
static bool f16_engine_on() {
    const int e = gemm_engine();
    return e == GGPM_ENGINE_AUTO || e == GGPM_ENGINE_TCGEN05_F16X3;
}

}  // namespace ggpm

using namespace ggpm;

extern "C" const char* ggpm_last_error(void) { return g_err; }
extern "C" int ggpm_abi_version(void) { return 1; }
extern "C" long long ggpm_launch_count(void) { return g_launches.load(); }
extern "C" int ggpm_get_gemm_engine(void) { return g_engine.load(); }
extern "C" int ggpm_set_gemm_engine(int engine) {
    GGPM_REQUIRE(engine >= GGPM_ENGINE_AUTO && engine <= GGPM_ENGINE_TCGEN05_F16X3, "unknown GEMM engine %d", engine);
    g_engine.store(engine);
    return 0;
}

extern "C" int ggpm_prof_enable(int on) {
    if (on && !g_prof_ev) {
        g_prof_ev = new cudaEvent_t[2 * kProfMax];
        for (int i = 0; i < 2 * kProfMax; ++i) GGPM_CUDA(cudaEventCreate(&g_prof_ev[i]));
    }
    g_prof_on = on != 0;
    if (on) g_prof_n = 0;
    return 0;
}

extern "C" int ggpm_prof_collect(double* ms, long long* count, double* work, int n_class) {
    GGPM_REQUIRE(ms && count && work && n_class >= PROF_NCLASS, "prof_collect: need %d classes", PROF_NCLASS);
    for (int c = 0; c < n_class; ++c) { ms[c] = 0; count[c] = 0; work[c] = 0; }
    GGPM_CUDA(cudaDeviceSynchronize());
    for (int i = 0; i < g_prof_n; ++i) {
        float t = 0.f;
        GGPM_CUDA(cudaEventElapsedTime(&t, g_prof_ev[2 * i], g_prof_ev[2 * i + 1]));
        ms[g_prof_cls[i]] += t;
        count[g_prof_cls[i]] += 1;
        work[g_prof_cls[i]] += g_prof_work[i];
    }
    g_prof_n = 0;
    return 0;
}

extern "C" int ggpm_gemm(const float* A, long long sam, long long sak, const float* B, long long sbk, long long sbn,
                         float* C, int ldc, int M, int N, int K, float beta, const float* bias, int act, int zero_row0,
                         int split_k, void* stream) {
    GGPM_REQUIRE(A && B && C, "gemm: null pointer");
    GemmArgs g;
    g.A = A; g.sam = sam; g.sak = sak; g.B = B; g.sbk = sbk; g.sbn = sbn; g.C = C; g.ldc = ldc;
    g.M = M; g.N = N; g.K = K; g.beta = beta; g.bias = bias; g.act = act; g.zero_row0 = zero_row0; g.split_k = split_k;
    return launch_gemm_nt(g, (cudaStream_t)stream);
}

extern "C" int ggpm_gemm_fused(const float* A, long long sam, long long sak, const float* B, long long sbk, long long sbn,
                               float* C, int ldc, int M, int N, int K, float beta, const float* bias, int act,
                               int zero_row0, int split_k, const int32_t* a_row_ptr, const int32_t* a_col,
                               const int32_t* b_row_ptr, const int32_t* b_col, int split_dim, int split_at,
                               const float* A2, long long lda2, int tile_n, const float* B_hi, const float* B_lo,
                               void* stream) {
    GGPM_REQUIRE(A && B && C, "gemm_fused: null pointer");
    GGPM_REQUIRE((a_row_ptr == nullptr) == (a_col == nullptr) && (b_row_ptr == nullptr) == (b_col == nullptr),
                 "gemm_fused: CSR arrays come in pairs");
    GGPM_REQUIRE(split_dim >= 0 && split_dim <= 3, "gemm_fused: split_dim must be 0..3");
    GemmArgs g;
    g.A = A; g.sam = sam; g.sak = sak; g.B = B; g.sbk = sbk; g.sbn = sbn; g.C = C; g.ldc = ldc;
    g.M = M; g.N = N; g.K = K; g.beta = beta; g.bias = bias; g.act = act; g.zero_row0 = zero_row0; g.split_k = split_k;
    g.a_row_ptr = a_row_ptr; g.a_col = a_col; g.b_row_ptr = b_row_ptr; g.b_col = b_col;
    g.split_dim = split_dim; g.split_at = split_at; g.A2 = A2; g.lda2 = lda2; g.tile_n = tile_n;
    g.B_hi = B_hi; g.B_lo = B_lo;
    GGPM_REQUIRE(gemm_tc_eligible(g), "gemm_fused: needs the tcgen05 engine and 16-byte aligned, 4-float strided operands");
    return launch_gemm_tc(g, (cudaStream_t)stream);
}

extern "C" int ggpm_f16_prepare_weights(const float* W, int rows, int cols, int transpose, void* hi, void* lo, float* scal,
                                        void* stream) {
    GGPM_REQUIRE(W && hi && lo && scal && rows > 0 && cols > 0, "f16_prepare_weights: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    GGPM_CUDA(cudaMemsetAsync(scal, 0, 2 * sizeof(float), st));
    GGPM_TRY(launch_absmax(W, (long long)rows * cols, scal, st));
    if (transpose) return launch_f16_split_transpose(W, rows, cols, scal, hi, lo, scal + 1, st);
    return launch_f16_split(W, (long long)rows * cols, scal, hi, lo, scal + 1, st);
}

extern "C" int ggpm_gemm_f16x3(const float* A, long long lda, const void* B_hi, const void* B_lo, const float* b_scale,
                               long long ldb, float* C, int ldc, int M, int N, int K, float beta, const float* bias,
                               int act, int zero_row0, const int32_t* a_row_ptr, const int32_t* a_col, int split_dim,
                               int split_at, const float* A2, long long lda2, const float* a_amax, int a_amax_n,
                               float a_amax_const, const int32_t* a_deg, int tile_n, void* stream) {
    GGPM_REQUIRE(A && B_hi && B_lo && b_scale && C, "gemm_f16x3: null pointer");
    GGPM_REQUIRE((a_row_ptr == nullptr) == (a_col == nullptr), "gemm_f16x3: CSR arrays come in pairs");
    GGPM_REQUIRE(split_dim >= 0 && split_dim <= 2, "gemm_f16x3: split_dim must be 0..2");
    GemmArgs g;
    g.A = A; g.sam = lda; g.sak = 1; g.B = nullptr; g.sbk = 1; g.sbn = ldb; g.C = C; g.ldc = ldc;
    g.M = M; g.N = N; g.K = K; g.beta = beta; g.bias = bias; g.act = act; g.zero_row0 = zero_row0;
    g.a_row_ptr = a_row_ptr; g.a_col = a_col; g.split_dim = split_dim; g.split_at = split_at; g.A2 = A2; g.lda2 = lda2;
    g.tile_n = tile_n; g.Bh_hi = B_hi; g.Bh_lo = B_lo; g.b_scale = b_scale;
    g.a_amax = a_amax; g.a_amax_n = a_amax_n; g.a_amax_const = a_amax_const; g.a_deg = a_deg;
    GGPM_REQUIRE(gemm_f16_eligible(g), "gemm_f16x3: needs the f16x3 engine, 16-byte aligned operands, K %% 4 == 0, ldb %% 8 == 0");
    return launch_gemm_f16(g, (cudaStream_t)stream);
}

extern "C" int ggpm_gemm_f16x3_planes(const void* A_hi, const void* A_lo, const float* a_scale, long long lda,
                                      const void* B_hi, const void* B_lo, const float* b_scale, long long ldb, float* C,
                                      int ldc, int M, int N, int K, float beta, int tn, void* stream) {
    GGPM_REQUIRE(A_hi && A_lo && a_scale && B_hi && B_lo && b_scale && C, "gemm_f16x3_planes: null pointer");
    GGPM_REQUIRE(beta == 0.f || beta == 1.f, "gemm_f16x3_planes: beta must be 0 or 1");
    GemmArgs g;
    g.Ah_hi = A_hi; g.Ah_lo = A_lo; g.a_scale = a_scale; g.lda_h = lda; g.f16_tn = tn ? 1 : 0;
    g.Bh_hi = B_hi; g.Bh_lo = B_lo; g.b_scale = b_scale; g.sbn = ldb;
    g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K; g.beta = beta;
    GGPM_REQUIRE(gemm_f16_eligible(g), "gemm_f16x3_planes: needs the f16x3 engine, 16-byte aligned planes, strides %% 8 == 0");
    return launch_gemm_f16(g, (cudaStream_t)stream);
}

extern "C" int ggpm_tf32_split(const float* w, long long n, float* hi, float* lo, void* stream) {
    GGPM_REQUIRE(w && hi && lo && n >= 0, "tf32_split: null pointer");
    return launch_tf32_split(w, n, hi, lo, (cudaStream_t)stream);
}

// ----------------------------------------------------------------------------------------------------
// LSTM
// ----------------------------------------------------------------------------------------------------
static int check_dims(const char* who, int E1, int H, int Hp, int depth) {
    GGPM_REQUIRE(E1 >= 1 && H >= 1 && depth >= 1, "%s: bad sizes E1=%d H=%d depth=%d", who, E1, H, depth);
    GGPM_REQUIRE(Hp >= H && Hp % 32 == 0, "%s: Hp=%d must be a multiple of 32 and >= H=%d", who, Hp, H);
    return 0;
}

// which optional buffers the LSTM layer calls will use under the current engine (so the caller can size them):
// bit 0: f16x3 engine for the gate / state-gradient GEMMs; bit 1: fp16 h planes + Z planes (TMA-fed backward GEMMs),
// in which case hist_h may be a 2-slot rotating buffer (h_rotate = 1)
extern "C" int ggpm_lstm_plan(int Hp, int depth) {
    const int e = gemm_engine();
    if (e == GGPM_ENGINE_SIMT_FP32 || Hp % 128 != 0 || depth <= 1) return 0;
    if (!f16_engine_on() || !f16_scalars_fit(Hp, depth)) return 0;
    return 3;
}

extern "C" int ggpm_lstm_fwd(int E1, int H, int Hp, int depth, const int32_t* b_ptr, const int32_t* b_col,
                             const float* Yn, const int32_t* src, const float* T, const int32_t* k1, const int32_t* k2,
                             const float* Ug, const float* Uf, float* hist_h, float* hist_c, float* hist_gp,
                             float* s_scratch, float* w_split, void* hist_h16, int keep_history, int h_rotate,
                             void* stream) {
    GGPM_TRY(check_dims("lstm_fwd", E1, H, Hp, depth));
    GGPM_REQUIRE(b_ptr && b_col && Yn && src && T && k1 && Ug && Uf && hist_h && hist_c && hist_gp && s_scratch,
                 "lstm_fwd: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t row = (size_t)E1 * Hp;
    const XTabArgs x{Yn, src, T, k1, k2};
    const bool tc_ok = tc_layer_ok(Hp, {hist_h, hist_c, hist_gp, Ug, Uf});
    const bool fused_w = tc_ok && Uf == Ug + (size_t)3 * Hp * Hp && Hp % 128 == 0;
    // weights are split into tf32 hi / lo planes once per call and then streamed by TMA in every round
    const size_t wn = (size_t)4 * Hp * Hp;
    const bool w_tma = fused_w && w_split != nullptr && ((uintptr_t)w_split & 15) == 0 && depth > 1;
    // f16x3 engine: fp16 hi / lo planes of the scaled weights in the first half of w_split, scalars behind them
    const bool f16 = w_tma && f16_engine_on() && f16_scalars_fit(Hp, depth);
    F16Work fw = f16_work(w_split, Hp);
    if (f16) {
        GGPM_CUDA(cudaMemsetAsync(fw.scal, 0, sizeof(float) * 16, st));
        GGPM_TRY(launch_absmax(Ug, (long long)wn, fw.scal + 0, st));
        GGPM_TRY(launch_f16_split(Ug, (long long)wn, fw.scal + 0, fw.hi, fw.lo, fw.scal + 1, st));
        GGPM_TRY(launch_max_degree(b_ptr, E1, reinterpret_cast<int*>(fw.scal + 2), st));
        static const float kH16 = 8192.f;              // GGPM_H16_SCALE of the forward cell kernel, as a device scalar
        GGPM_CUDA(cudaMemcpyAsync(fw.scal + 6, &kH16, sizeof(float), cudaMemcpyHostToDevice, st));
    } else if (w_tma) {
        GGPM_TRY(launch_tf32_split(Ug, (long long)wn, w_split, w_split + wn, st));
    }
    const bool h16 = f16 && keep_history && hist_h16 != nullptr && ((uintptr_t)hist_h16 & 15) == 0;
    GGPM_REQUIRE(!h_rotate || !keep_history || h16, "lstm_fwd: a rotating hist_h needs the fp16 h planes (ggpm_lstm_plan bit 1)");
    // large layers (the plane-fed path below) read h^{t-1} only through its fp16 planes: the fp32 copy is written by the
    // last round alone (it is what the call returns)
    const bool planes_only = f16 && h16 && E1 >= sgather_min_rows() && Hp % 64 == 0 && ((uintptr_t)s_scratch & 15) == 0 &&
                             fused_w && tc_ok;
    for (int t = 1; t <= depth; ++t) {
        const int cur = keep_history ? (t - 1) : ((t - 1) & 1);
        const int prev = keep_history ? (t - 2) : ((t - 2) & 1);
        float* h_cur = hist_h + (h_rotate ? (cur & 1) : cur) * row;
        float* c_cur = hist_c + cur * row;
        float* gp = hist_gp + (size_t)cur * 4 * row;
        // fp16 planes of h^t (2 bytes per element): [cur][hi | lo][E1 * Hp]
        char* h16_hi = h16 ? (char*)hist_h16 + (size_t)cur * 4 * row : nullptr;
        char* h16_lo = h16 ? h16_hi + 2 * row : nullptr;
        float* h_out = (planes_only && t < depth) ? nullptr : h_cur;
        if (t == 1) {
            GGPM_TRY(launch_lstm_cell_fwd(true, E1, H, Hp, b_ptr, b_col, x, nullptr, nullptr, h_out, c_cur, h16_hi, h16_lo, st));
            continue;
        }
        const float* h_prev = hist_h + (h_rotate ? (prev & 1) : prev) * row;
        const float* c_prev = hist_c + prev * row;
        // S = sum of neighbour states; [G_i|G_o|G_u] = S Ug^T ; P = h_prev Uf^T
        if (tc_ok) {
            // tcgen05: the gather-sum is the A-operand prologue of the gate GEMM; S is never written
            GemmArgs g;
            g.A = h_prev; g.sam = Hp; g.sak = 1; g.B = Ug; g.sbk = 1; g.sbn = Hp; g.C = gp; g.ldc = 4 * Hp;
            g.M = E1; g.K = Hp; g.a_row_ptr = b_ptr; g.a_col = b_col;
            if (fused_w) {   // one launch: columns >= 3Hp (the forget projection) read h_prev un-gathered
                g.N = 4 * Hp; g.split_dim = 1; g.split_at = 3 * Hp;
                if (planes_only) {
                    // large layer: S as fp16 planes from one gather kernel (s_scratch holds both planes), then two GEMMs
                    // whose operands all arrive by TMA: [G_i|G_o|G_u] = S Ug^T and P = h^{t-1} Uf^T from the saved h planes
                    char* s_hi = (char*)s_scratch;
                    char* s_lo = s_hi + 2 * row;
                    float* s_scale = fw.scal + 5;
                    const char* hp_hi = (const char*)hist_h16 + (size_t)prev * 4 * row;
                    GGPM_TRY(launch_lstm_sgather16(E1, Hp, b_ptr, b_col, hp_hi, hp_hi + 2 * row,
                                                   reinterpret_cast<const int*>(fw.scal + 2), s_hi, s_lo, s_scale, st));
                    GemmArgs a;
                    a.Ah_hi = s_hi; a.Ah_lo = s_lo; a.lda_h = Hp; a.a_scale = s_scale;
                    a.Bh_hi = fw.hi; a.Bh_lo = fw.lo; a.sbn = Hp; a.b_scale = fw.scal + 1;
                    a.C = gp; a.ldc = 4 * Hp; a.M = E1; a.N = 3 * Hp; a.K = Hp;
                    if (fused_gates()) {
                        a.act = GGPM_ACT_LSTM_GATES;
                        a.gx_Yn = Yn; a.gx_src = src; a.gx_T = T; a.gx_k1 = k1; a.gx_k2 = k2; a.gx_ld = 4 * Hp; a.gx_hp = Hp;
                    }
                    GGPM_TRY(launch_gemm_f16(a, st));
                    GemmArgs q;
                    q.Ah_hi = hp_hi; q.Ah_lo = hp_hi + 2 * row; q.lda_h = Hp; q.a_scale = fw.scal + 6;      // = 2^13
                    q.Bh_hi = (const char*)fw.hi + (size_t)3 * Hp * Hp * 2; q.Bh_lo = (const char*)fw.lo + (size_t)3 * Hp * Hp * 2;
                    q.sbn = Hp; q.b_scale = fw.scal + 1;
                    q.C = gp + 3 * Hp; q.ldc = 4 * Hp; q.M = E1; q.N = Hp; q.K = Hp;
                    GGPM_TRY(launch_gemm_f16(q, st));
                } else if (f16) {   // |h| <= 1, so every gathered row is bounded by the largest neighbour count
                    g.Bh_hi = fw.hi; g.Bh_lo = fw.lo; g.b_scale = fw.scal + 1;
                    g.a_amax_const = 1.f; g.a_deg = reinterpret_cast<const int*>(fw.scal + 2);
                    GGPM_TRY(launch_gemm_f16(g, st));
                } else {
                    if (w_tma) { g.B_hi = w_split; g.B_lo = w_split + wn; }
                    GGPM_TRY(launch_gemm_tc(g, st));
                }
            } else {
                g.N = 3 * Hp;
                GGPM_TRY(launch_gemm_tc(g, st));
                GGPM_TRY(gemm_nt(h_prev, Hp, Uf, Hp, gp + 3 * Hp, 4 * Hp, E1, Hp, Hp, 0.f, nullptr, GGPM_ACT_NONE, 0, st));
            }
        } else {
            GGPM_TRY(launch_gather_sum(h_prev, Hp, b_ptr, b_col, s_scratch, Hp, E1, Hp, false, st));
            GGPM_TRY(gemm_nt(s_scratch, Hp, Ug, Hp, gp, 4 * Hp, E1, 3 * Hp, Hp, 0.f, nullptr, GGPM_ACT_NONE, 0, st));
            GGPM_TRY(gemm_nt(h_prev, Hp, Uf, Hp, gp + 3 * Hp, 4 * Hp, E1, Hp, Hp, 0.f, nullptr, GGPM_ACT_NONE, 0, st));
        }
        GGPM_TRY(launch_lstm_cell_fwd(false, E1, H, Hp, b_ptr, b_col, x, gp, c_prev, h_out, c_cur, h16_hi, h16_lo, st,
                                      planes_only && fused_gates()));
    }
    return 0;
}

extern "C" int ggpm_lstm_bwd(int E1, int H, int Hp, int depth, const int32_t* b_ptr, const int32_t* b_col,
                             const int32_t* bt_ptr, const int32_t* bt_row, const float* Yn, const int32_t* src,
                             const float* T, const int32_t* k1, const int32_t* k2, const float* Ug, const float* Uf,
                             const float* hist_h, float* hist_c, float* hist_gp, float* dH_final,
                             const float* dC_final, float* dX, float* dUg, float* dUf, float* dA_unused, float* dC, float* dP,
                             float* dS, float* S, float* dHn, float* w_split, const void* hist_h16, void* z_planes,
                             int h_rotate, const int32_t* bt_pos, float* pair_buf, long long pair_rows, void* stream) {
    GGPM_TRY(check_dims("lstm_bwd", E1, H, Hp, depth));
    GGPM_REQUIRE(b_ptr && b_col && bt_ptr && bt_row && Yn && src && T && k1 && Ug && Uf && hist_h && hist_c && hist_gp &&
                     dH_final && dX && dUg && dUf && dC && dP && dS && S && dHn,
                 "lstm_bwd: null pointer");
    (void)dA_unused;   // round-1 ABI scratch: the gate gradients of a round now overwrite that round's slot of hist_gp
    cudaStream_t st = (cudaStream_t)stream;
    const size_t row = (size_t)E1 * Hp;
    const XTabArgs x{Yn, src, T, k1, k2};
    float* dH = dH_final;
    float* dH_next = dHn;
    const float* dC_cur = dC_final;  // dL/dc^depth (NULL: zero — the encoders consume only h)
    GGPM_REQUIRE(((uintptr_t)dC_final & 15) == 0, "lstm_bwd: dC_final must be 16-byte aligned");
    const bool tc_ok = tc_layer_ok(Hp, {hist_h, hist_c, hist_gp, Ug, Uf, dUg, dUf, dP, dH_final, dHn});
    const bool fused_w = tc_ok && Uf == Ug + (size_t)3 * Hp * Hp && dUf == dUg + (size_t)3 * Hp * Hp && Hp % 128 == 0;
    // state gradient dH = [..] [Ug; Uf]: the transposed weights [Hp, 4Hp] are split into tf32 planes once, then TMA-fed
    const size_t wn = (size_t)4 * Hp * Hp;
    const bool w_tma = fused_w && w_split != nullptr && ((uintptr_t)w_split & 15) == 0 && depth > 1;
    const bool f16 = w_tma && f16_engine_on() && f16_scalars_fit(Hp, depth);
    F16Work fw = f16_work(w_split, Hp);
    if (f16) {
        GGPM_CUDA(cudaMemsetAsync(fw.scal, 0, sizeof(float) * (16 + 16 * (size_t)(depth + 1)), st));
        GGPM_TRY(launch_absmax(Ug, (long long)wn, fw.scal + 0, st));
        GGPM_TRY(launch_f16_split_transpose(Ug, 4 * Hp, Hp, fw.scal + 0, fw.hi, fw.lo, fw.scal + 1, st));
        GGPM_TRY(launch_max_degree(bt_ptr, E1, reinterpret_cast<int*>(fw.scal + 2), st));
    } else if (w_tma) {
        GGPM_TRY(launch_tf32_split_transpose(Ug, 4 * Hp, Hp, w_split, w_split + wn, st));
    }
    // fully TMA-fed backward GEMMs: Z planes (written by the zgather kernel) and the h planes saved by the forward
    const bool zmode = f16 && hist_h16 != nullptr && z_planes != nullptr && Hp % 64 == 0 &&
                       (((uintptr_t)hist_h16 | (uintptr_t)z_planes) & 15) == 0;
    GGPM_REQUIRE(zmode || !h_rotate, "lstm_bwd: the forward kept only a rotating hist_h, which needs the fp16-plane path");
    char* z_hi = (char*)z_planes;
    char* z_lo = z_hi + (size_t)8 * row;                  // 4 * row halves
    // pair mode (plane path only): the two per-pair terms of a round are stored in pair order and gathered over the
    // reverse CSR where they are consumed, instead of being reduced into zeroed tables with atomics
    // pair_buf: [3, pair_rows, Hp] = daf pairs | dchat * f pairs of even rounds | of odd rounds; pair_rows >= nnz
    // the forward kept ACTIVATED gates in hist_gp on the plane-fed path when GGPM_FUSED_GATES is on (same predicate as there)
    const bool gates_act = fused_gates() && zmode && fused_w && tc_ok && E1 >= sgather_min_rows() && Hp % 64 == 0;
    const bool pairs = zmode && bt_pos != nullptr && pair_buf != nullptr && pair_rows > 0 && ((uintptr_t)pair_buf & 15) == 0;
    const size_t pair_slot = (size_t)(pair_rows > 0 ? pair_rows : 0) * Hp;
    SideStream* side = zmode ? side_stream() : nullptr;
    bool side_busy = false;                               // a weight-gradient GEMM may still be reading the Z planes
    for (int t = depth; t >= 1; --t) {
        float* c_cur = hist_c + (size_t)(t - 1) * row;          // consumed: receives sum_n daf_en of the round
        float* gp = hist_gp + (size_t)(t - 1) * 4 * row;        // consumed: [G_i|G_o|G_u] become [dai|dao|dau]
        float* dA = gp;
        float* pairA = pairs ? pair_buf : nullptr;
        float* pairB = pairs ? pair_buf + (size_t)(1 + (t & 1)) * pair_slot : nullptr;
        const float* pairB_in = (pairs && t < depth) ? pair_buf + (size_t)(1 + ((t + 1) & 1)) * pair_slot : nullptr;
        const float* dC_dense = pairs ? (t == depth ? dC_final : nullptr) : dC_cur;
        if (t == 1) {
            GGPM_TRY(launch_lstm_cell_bwd(true, E1, H, Hp, b_ptr, b_col, x, gp, nullptr, c_cur, dH, dC_dense, nullptr, nullptr,
                                          nullptr, st, bt_ptr, bt_pos, pairB_in, pairA, pairB, pairs));
            break;
        }
        const float* h_prev = h_rotate ? nullptr : hist_h + (size_t)(t - 2) * row;     // only read without the plane path
        const float* c_prev = hist_c + (size_t)(t - 2) * row;
        float* dC_prev = dC + (size_t)(t & 1) * row;
        if (!pairs) GGPM_TRY(launch_zero2(dP, dC_prev, row, st));
        // the cell kernel also records max |gate-gradient term| of this round (16 slots): the scale of the f16x3 GEMM
        float* amax_t = f16 ? fw.scal + 16 + 16 * (size_t)t : nullptr;
        GGPM_TRY(launch_lstm_cell_bwd(false, E1, H, Hp, b_ptr, b_col, x, gp, c_prev, c_cur, dH, dC_dense, dP, dC_prev, amax_t, st,
                                      bt_ptr, bt_pos, pairB_in, pairA, pairB, pairs, gates_act));
        if (zmode) {
            // Z = [gather^T(dA[:, :3Hp]) | dP] as scaled fp16 planes; then dH^{t-1} = Z [Ug; Uf] and
            // [dUg; dUf] += Z^T h^{t-1}: two plain GEMMs whose operands all arrive by TMA
            float* z_scale = fw.scal + 4;
            if (side_busy) GGPM_CUDA(cudaStreamWaitEvent(st, side->join, 0));     // Z planes / scale are rewritten next
            GGPM_TRY(launch_lstm_zgather(E1, Hp, bt_ptr, bt_row, dA, dP, bt_pos, pairA, amax_t,
                                         reinterpret_cast<const int*>(fw.scal + 2), z_hi, z_lo, z_scale, st));
            GemmArgs w;
            w.f16_tn = 1; w.Ah_hi = z_hi; w.Ah_lo = z_lo; w.lda_h = 4 * Hp; w.a_scale = z_scale;
            w.Bh_hi = (const char*)hist_h16 + (size_t)(t - 2) * 4 * row; w.Bh_lo = (const char*)w.Bh_hi + 2 * row;
            w.sbn = Hp; w.b_scale_const = 8192.f;          // GGPM_H16_SCALE of the forward cell kernel
            w.C = dUg; w.ldc = Hp; w.M = 4 * Hp; w.N = Hp; w.K = E1; w.beta = 1.f;
            if (side == nullptr) GGPM_TRY(launch_gemm_f16(w, st));
            GemmArgs d;
            d.Ah_hi = z_hi; d.Ah_lo = z_lo; d.lda_h = 4 * Hp; d.a_scale = z_scale;
            d.Bh_hi = fw.hi; d.Bh_lo = fw.lo; d.sbn = 4 * Hp; d.b_scale = fw.scal + 1;
            d.C = dH_next; d.ldc = Hp; d.M = E1; d.N = Hp; d.K = 4 * Hp;
            if (side != nullptr) {
                // critical path first (its CTAs are dispatched ahead of the weight-gradient GEMM's), then the fork
                GGPM_CUDA(cudaEventRecord(side->fork, st));                        // Z planes of this round complete
                GGPM_TRY(launch_gemm_f16(d, st));
                GGPM_CUDA(cudaStreamWaitEvent(side->s, side->fork, 0));
                GGPM_TRY(launch_gemm_f16(w, side->s));
                GGPM_CUDA(cudaEventRecord(side->join, side->s));
                side_busy = true;
            } else {
                GGPM_TRY(launch_gemm_f16(d, st));
            }
        } else if (tc_ok) {
            // weight gradients: dUg += dA[:, :3Hp]^T S^t (S re-gathered inside the B-operand prologue) ;
            // dUf += dP^T h^{t-1}
            GemmArgs w;
            w.A = dA; w.sam = 1; w.sak = 4 * Hp; w.B = h_prev; w.sbk = Hp; w.sbn = 1; w.C = dUg; w.ldc = Hp;
            w.N = Hp; w.K = E1; w.beta = 1.f; w.b_row_ptr = b_ptr; w.b_col = b_col;
            if (fused_w) {   // rows >= 3Hp of [dUg; dUf]: A = dP, B = h_prev un-gathered
                w.M = 4 * Hp; w.split_dim = 3; w.split_at = 3 * Hp; w.A2 = dP; w.lda2 = Hp;
                w.split_k = wgrad_splits(4 * Hp, Hp, E1);
                GGPM_TRY(launch_gemm_tc(w, st));
            } else {
                w.M = 3 * Hp; w.split_k = wgrad_splits(3 * Hp, Hp, E1);
                GGPM_TRY(launch_gemm_tc(w, st));
                GGPM_TRY(gemm_tn_acc(dP, Hp, h_prev, Hp, dUf, Hp, Hp, Hp, E1, st));
            }
            // state gradient: dH^{t-1}[n] = sum_{e: n in N(e)} dA[e, :3Hp] Ug + dP[n] Uf
            //               = [gather^T(dA[:, :3Hp]) | dP] [Ug; Uf]   (gather moved in front of the GEMM by linearity)
            GemmArgs d;
            d.A = dA; d.sam = 4 * Hp; d.sak = 1; d.B = Ug; d.sbk = Hp; d.sbn = 1; d.C = dH_next; d.ldc = Hp;
            d.M = E1; d.N = Hp; d.a_row_ptr = bt_ptr; d.a_col = bt_row;
            if (fused_w) {
                d.K = 4 * Hp; d.split_dim = 2; d.split_at = 3 * Hp; d.A2 = dP; d.lda2 = Hp;
                if (f16) {     // rows are sums of <= (largest reverse-neighbour count) terms, each <= amax_t in magnitude
                    d.B = Ug; d.sbk = 1; d.sbn = 4 * Hp; d.Bh_hi = fw.hi; d.Bh_lo = fw.lo; d.b_scale = fw.scal + 1;
                    d.a_amax = amax_t; d.a_amax_n = 16; d.a_deg = reinterpret_cast<const int*>(fw.scal + 2);
                    GGPM_TRY(launch_gemm_f16(d, st));
                } else {
                    if (w_tma) {   // B(k, n) = W^T[n][k]: k-contiguous planes
                        d.B = w_split; d.sbk = 1; d.sbn = 4 * Hp; d.B_hi = w_split; d.B_lo = w_split + wn;
                    }
                    GGPM_TRY(launch_gemm_tc(d, st));
                }
            } else {
                d.K = 3 * Hp;
                GGPM_TRY(launch_gemm_tc(d, st));
                GGPM_TRY(gemm_nn(dP, Hp, Uf, Hp, dH_next, Hp, E1, Hp, Hp, 1.f, st));
            }
        } else {
            // weight gradients: dUg += dA[:, :3Hp]^T S^t ; dUf += dP^T h^{t-1}
            GGPM_TRY(launch_gather_sum(h_prev, Hp, b_ptr, b_col, S, Hp, E1, Hp, false, st));
            GGPM_TRY(gemm_tn_acc(dA, 4 * Hp, S, Hp, dUg, Hp, 3 * Hp, Hp, E1, st));
            GGPM_TRY(gemm_tn_acc(dP, Hp, h_prev, Hp, dUf, Hp, Hp, Hp, E1, st));
            // state gradients: dS = dA[:, :3Hp] Ug ; dH^{t-1}[n] = sum_{e: n in N(e)} dS[e] + (dP Uf)[n]
            GGPM_TRY(gemm_nn(dA, 4 * Hp, Ug, Hp, dS, Hp, E1, Hp, 3 * Hp, 0.f, st));
            GGPM_TRY(launch_gather_sum(dS, Hp, bt_ptr, bt_row, dH_next, Hp, E1, Hp, false, st));
            GGPM_TRY(gemm_nn(dP, Hp, Uf, Hp, dH_next, Hp, E1, Hp, Hp, 1.f, st));
        }
        float* tmp = dH; dH = dH_next; dH_next = tmp;
        dC_cur = dC_prev;
    }
    if (side_busy) GGPM_CUDA(cudaStreamWaitEvent(st, side->join, 0));   // dUg / dUf complete for the caller's stream
    // gradient of the hoisted x-projection: the per-round gate gradients, summed over the rounds in one streaming pass
    GGPM_TRY(launch_lstm_dx_reduce(E1, Hp, depth, hist_gp, hist_c, dX, st));
    return 0;
}

// ----------------------------------------------------------------------------------------------------
// GRU      aux row layout (5*Hp): [S | Gz | Gh | Pr | Sg]
// ----------------------------------------------------------------------------------------------------
extern "C" int ggpm_gru_fwd(int E1, int H, int Hp, int depth, const int32_t* b_ptr, const int32_t* b_col,
                            const float* Yn, const int32_t* src, const float* T, const int32_t* k1, const int32_t* k2,
                            const float* Uz, const float* Ur, const float* br, const float* Uh, float* hist_h,
                            float* hist_aux, int keep_history, void* stream) {
    GGPM_TRY(check_dims("gru_fwd", E1, H, Hp, depth));
    GGPM_REQUIRE(b_ptr && b_col && Yn && src && T && k1 && Uz && Ur && br && Uh && hist_h && hist_aux, "gru_fwd: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t row = (size_t)E1 * Hp;
    const XTabArgs x{Yn, src, T, k1, k2};
    GGPM_CUDA(cudaMemsetAsync(hist_h, 0, sizeof(float) * row, st));   // h^0 = 0
    for (int t = 1; t <= depth; ++t) {
        const float* h_prev = hist_h + (size_t)(keep_history ? (t - 1) : ((t - 1) & 1)) * row;
        float* h_cur = hist_h + (size_t)(keep_history ? t : (t & 1)) * row;
        float* aux = hist_aux + (size_t)(keep_history ? (t - 1) : 0) * 5 * row;
        GGPM_TRY(launch_gather_sum(h_prev, Hp, b_ptr, b_col, aux, 5 * Hp, E1, Hp, false, st));                          // S
        GGPM_TRY(gemm_nt(aux, 5 * Hp, Uz, Hp, aux + Hp, 5 * Hp, E1, Hp, Hp, 0.f, nullptr, GGPM_ACT_NONE, 0, st));        // Gz
        GGPM_TRY(gemm_nt(h_prev, Hp, Ur, Hp, aux + 3 * Hp, 5 * Hp, E1, Hp, Hp, 0.f, br, GGPM_ACT_NONE, 0, st));          // Pr
        GGPM_TRY(launch_gru_gate(E1, Hp, b_ptr, b_col, x, h_prev, aux, st));                                            // Sg
        GGPM_TRY(gemm_nt(aux + 4 * Hp, 5 * Hp, Uh, Hp, aux + 2 * Hp, 5 * Hp, E1, Hp, Hp, 0.f, nullptr, GGPM_ACT_NONE, 0, st));  // Gh
        GGPM_TRY(launch_gru_out(E1, H, Hp, x, aux, h_cur, st));
    }
    return 0;
}

extern "C" int ggpm_gru_bwd(int E1, int H, int Hp, int depth, const int32_t* b_ptr, const int32_t* b_col,
                            const float* Yn, const int32_t* src, const float* T, const int32_t* k1, const int32_t* k2,
                            const float* Uz, const float* Ur, const float* Uh, const float* hist_h, const float* hist_aux,
                            float* dH_final, float* dX, float* dUz, float* dUr, float* dbr, float* dUh, float* dA,
                            float* dS, float* dSg, float* dPr, float* dHn, void* stream) {
    GGPM_TRY(check_dims("gru_bwd", E1, H, Hp, depth));
    GGPM_REQUIRE(b_ptr && b_col && Yn && src && T && k1 && Uz && Ur && Uh && hist_h && hist_aux && dH_final && dX && dUz &&
                     dUr && dbr && dUh && dA && dS && dSg && dPr && dHn,
                 "gru_bwd: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t row = (size_t)E1 * Hp;
    const XTabArgs x{Yn, src, T, k1, k2};
    float* dH = dH_final;
    float* dH_next = dHn;
    for (int t = depth; t >= 1; --t) {
        const float* h_prev = hist_h + (size_t)(t - 1) * row;
        const float* aux = hist_aux + (size_t)(t - 1) * 5 * row;
        const bool overwrite = (t == depth);
        GGPM_TRY(launch_gru_out_bwd(E1, H, Hp, x, aux, dH, dA, dS, st));        // dA = [daz|dah], dS = dH (1-z)
        if (t == 1) {   // h^0 = 0: no neighbour terms, no weight gradients from this round
            GGPM_TRY(launch_gru_pair_bwd(E1, Hp, b_ptr, b_col, x, aux, h_prev, dA, dS, dSg, dX, overwrite, false, dPr, dH_next, st));
            break;
        }
        GGPM_TRY(gemm_nn(dA, 2 * Hp, Uz, Hp, dS, Hp, E1, Hp, Hp, 1.f, st));       // dS += daz Uz
        GGPM_TRY(gemm_nn(dA + Hp, 2 * Hp, Uh, Hp, dSg, Hp, E1, Hp, Hp, 0.f, st)); // dSg = dah Uh
        GGPM_TRY(gemm_tn_acc(dA, 2 * Hp, aux, 5 * Hp, dUz, Hp, Hp, Hp, E1, st));                  // dUz += daz^T S
        GGPM_TRY(gemm_tn_acc(dA + Hp, 2 * Hp, aux + 4 * Hp, 5 * Hp, dUh, Hp, Hp, Hp, E1, st));    // dUh += dah^T Sg
        GGPM_CUDA(cudaMemsetAsync(dPr, 0, sizeof(float) * row, st));
        GGPM_CUDA(cudaMemsetAsync(dH_next, 0, sizeof(float) * row, st));
        GGPM_TRY(launch_gru_pair_bwd(E1, Hp, b_ptr, b_col, x, aux, h_prev, dA, dS, dSg, dX, overwrite, true, dPr, dH_next, st));
        GGPM_TRY(gemm_nn(dPr, Hp, Ur, Hp, dH_next, Hp, E1, Hp, Hp, 1.f, st));     // dH^{t-1} += dPr Ur
        GGPM_TRY(gemm_tn_acc(dPr, Hp, h_prev, Hp, dUr, Hp, Hp, Hp, E1, st));      // dUr += dPr^T h^{t-1}
        GGPM_TRY(ggpm_colsum(dPr, Hp, E1, Hp, dbr, 1, stream));                   // db_r += column sums of dPr
        float* tmp = dH; dH = dH_next; dH_next = tmp;
    }
    return 0;
}
