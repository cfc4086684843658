// qp_api.cu — host side of the C ABI declared in include/qpolicy.h.
//
// Owns device memory, the stream and the launch sequence; contains no arithmetic of the hot path
// (that is all in qp_kernels.cuh) and no CPU fallback: without a CUDA device qp_create fails.
#include "qp_kernels.cuh"
#include "qp_wide.cuh"
#include "qp_tc.cuh"
#include "qp_ep_tc.cuh"
#include "qp_ep_c2.cuh"
#include "qp_probe.cuh"
#include "qp_fanout.cuh"
#include "../../include/qpolicy.h"

#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cerrno>
#include <cstdlib>
#include <map>
#include <atomic>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <vector>

using namespace qp;

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}

#define QP_CUDA(expr)                                                                          \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return fail(QP_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));         \
    } while (0)

#define QP_TRY(expr)                   \
    do {                               \
        int _s = (expr);               \
        if (_s != QP_OK) return _s;    \
    } while (0)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;  // elements
    int ensure(size_t n) {
        if (n <= cap) return QP_OK;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        QP_CUDA(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)));
        cap = n;
        return QP_OK;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
    }
};

int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

// ---- NCCL, resolved at run time (dlopen) so that the library has no link-time dependency on it
// and shares the copy a host process may already have loaded (e.g. the one bundled with torch).
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
constexpr int kNcclFloat32 = 7, kNcclInt32 = 2, kNcclSum = 0, kNcclMin = 3;   // ncclDataType_t / ncclRedOp_t values of nccl.h

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
};
NcclApi g_nccl;

int nccl_load() {
    if (g_nccl.lib) return QP_OK;
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) return fail(QP_ENCCL, std::string("cannot load libnccl.so.2: ") + dlerror());
    NcclApi a;
    a.lib = lib;
    a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(lib, "ncclGetUniqueId");
    a.CommInitRank = (decltype(a.CommInitRank))dlsym(lib, "ncclCommInitRank");
    a.AllReduce = (decltype(a.AllReduce))dlsym(lib, "ncclAllReduce");
    a.CommDestroy = (decltype(a.CommDestroy))dlsym(lib, "ncclCommDestroy");
    a.GetErrorString = (decltype(a.GetErrorString))dlsym(lib, "ncclGetErrorString");
    a.GetVersion = (decltype(a.GetVersion))dlsym(lib, "ncclGetVersion");
    if (!a.GetUniqueId || !a.CommInitRank || !a.AllReduce || !a.CommDestroy || !a.GetErrorString)
        return fail(QP_ENCCL, "libnccl.so.2 lacks a required symbol");
    g_nccl = a;
    return QP_OK;
}

#define QP_NCCL(expr)                                                                             \
    do {                                                                                          \
        ncclResult_t _r = (expr);                                                                 \
        if (_r != 0) return fail(QP_ENCCL, std::string(#expr) + ": " + g_nccl.GetErrorString(_r)); \
    } while (0)

}  // namespace

namespace {
// Launch geometry and shared-memory budget of the register-tiled fp32 path for one shape.
struct Geometry {
    int step_warps = 0;                 // K1: one warp per 20 hidden columns
    size_t smem_step[3] = {0, 0, 0};
    int n_rt = 0, cg_max = 0, dw1_tiles = 0, dw1_threads = 0;   // K2
    size_t smem_dw1 = 0;
    bool big_ok = false;                // large-batch K2 (16 x 8 tiles, one split per SM) usable for this shape
    int big_n_rt = 0, big_n_ct = 0, big_threads = 0;
    size_t smem_big = 0;
    bool ep_ok = false;                 // persistent episode kernel usable for this shape
    int ep_n_rt = 0, ep_n_ct = 0, ep_threads = 0;
    size_t smem_ep = 0;
};

}  // namespace

struct qp_handle {
    qp_config cfg{};
    Hyper hp{};
    int64_t P = 0, P1 = 0;
    int device = 0, num_sms = 0, max_smem = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};

    DevBuf<float> params, acc, grad;
    // scratch shared by the episode and message paths
    DevBuf<float> dz1, xextra, part1, part2;
    int64_t Bpad = 0;
    // message path staging
    DevBuf<float> m_states, m_next, m_rewards, m_steps, m_q;
    DevBuf<int64_t> m_ids;
    DevBuf<int32_t> m_actions;
    // episodes
    bool has_env = false;
    DevBuf<float> prices;
    DevBuf<int32_t> series_of_env, shares;
    DevBuf<double> budget, share_value, portfolio, blocksum;
    DevBuf<int32_t> d_step;
    int64_t series_len = 0, n_series = 0, n_envs = 0, env_id_base = 0;
    double B0 = 0.0;
    int32_t N0 = 0;
    int64_t steps_done = 0;

    int64_t launches = 0;
    // data parallelism over GPUs: one SUM all-reduce of the gradient per env-step
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    // in-kernel all-reduce over NVLink peer memory (persistent kernel, world > 1)
    // exported to the other ranks: [256 B of flags | 2 * P floats] for episode_kernel, then, for episode_tc_kernel,
    // [2 x 8 x p_pad {value, tag} pairs] (p2p_tc_offset / p2p_pad below)
    void* p2p_block = nullptr;
    void* p2p_peer_block[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    bool p2p_ready = false, allow_p2p = true;
    bool p2p_external = false;          // the exchange blocks belong to the caller (qp_dist_p2p_attach): never freed / closed here
    void* p2p_mc = nullptr;             // multicast address of the blocks (NVLS): one store reaches every rank's copy
    unsigned int p2p_flag_base = 0;
    // batching front-end (config 5): child actors enqueue rows, one flush = one GPU launch per kind
    // batching front-end: two generations of queued messages per kind, in pinned host memory.  Actors reserve a slot
    // with one atomic add under a SHARED lock and copy their message straight into the staging buffer the flush
    // uploads from; the exclusive lock is taken only to grow a queue or to hand a generation to the flush.
    struct FrontQueue {
        float* states = nullptr; float* next = nullptr; float* scalars = nullptr; int64_t* ids = nullptr;
        int64_t cap = 0;                  // messages the buffers hold
        std::atomic<int64_t> count{0};    // slots handed out (may overshoot cap: those callers grow the queue and retry)
    };
    FrontQueue q_sel[2], q_upd[2];
    int cur_sel = 0, cur_upd = 0;       // generation the actors write to
    std::shared_mutex qrw;
    std::mutex flush_mu;                // flushes are serialised
    std::vector<int32_t> q_results;
    int64_t ticket_base = 0;            // ticket of q_results[0]
    int64_t sel_first_ticket = 0;       // ticket of slot 0 of the current select generation
    // UpdateQ counter and snapshot hook (QDecisionPolicyActor.scala:74, :76)
    int64_t iteration = 0;
    std::string snapshot_path;
    int64_t snapshot_every = 500;
    // persistent kernel with tensor-core contractions (qp_ep_tc.cuh)
    bool allow_ep_tc = true, ep_tc_ready = false;   // qp_set_option("ep_tc", 0) keeps the FFMA persistent kernel
    bool prices_exact = true;           // every price is a tf32 value (checked at qp_env_create)
    // router fan-out path (qp_fanout.cuh): episodes grouped by series, chunks of <= kFanChunk episodes of one series
    bool env_identity = false;          // episode e runs on series e
    DevBuf<float> c2_pT;                // pair kernel in waves: transposed price matrix [T][series padded to 64]
    CUtensorMap tm_c2_win{};
    bool c2_win_ready = false;
    bool allow_c2_waves = true;         // qp_set_option("ep_c2_waves"): the pair kernel also where a pair has several tiles
    int allow_fanout = 1;               // qp_set_option("fanout", 0 | 1 | 2): never / where it pays / wherever it applies
    bool fan_ok = false;
    int fan_n_chunks = 0;
    std::vector<int32_t> fan_chunks_host;
    DevBuf<int32_t> fan_chunks;
    DevBuf<float> fan_zw, fan_part;
    bool et_flow_ok = false;            // the shape takes phase B's bulk-copy branch (the only one with the NVLink exchange)
    // the same kernel on CTA pairs that split the hidden dimension (qp_ep_c2.cuh): the default where it applies
    bool allow_ep_c2 = true, ep_c2_ready = false;   // qp_set_option("ep_c2", 0) keeps episode_tc_kernel
    CUtensorMap tm_c2_hi, tm_c2_lo, tm_c2_part;
    const float* tm_c2_part_ptr = nullptr; int64_t tm_c2_part_pairs = 0;
    int c2_Hh = 0, c2_Nh = 0, c2_grid = 0, c2_dbg = 0;
    size_t c2_smem_bytes = 0;
    // several GPUs: what ALL ranks can do (MIN over the ranks at qp_env_create); a rank never picks a path alone
    bool agree_valid = false, agree_ep_tc = false, agree_fits = false, agree_fused = false, agree_p2p = false;
    bool agree_c2 = false, agree_rsag = false, agree_nvls = false;   // pair kernel, its exchange scheme, multicast stores
    DevBuf<int32_t> flags4;             // small device scratch (tf32-exactness flag, rank agreement)
    int32_t* host_flag = nullptr;       // pinned: the flag's copy back rides on a synchronisation that happens anyway
    DevBuf<float> et_w_hi, et_w_lo;
    CUtensorMap tm_et_hi, tm_et_lo, tm_et_part;
    const float* tm_et_part_ptr = nullptr; int64_t tm_et_part_tiles = 0;   // what tm_et_part was encoded for
    DevBuf<unsigned long long> et_published;    // split grid barrier of episode_tc_kernel
    int et_Kp = 0, et_nkb = 0, et_Np = 0, et_halves = 0;
    size_t et_smem = 0;
    Geometry geo;
    bool tile_ok = false;               // W1 fits in shared memory: per-tile kernels usable
    bool wide = false;                  // GEMM-organised path (qp_wide.cuh)
    DevBuf<float> wZ1, wdZ1, wdz2;
    DevBuf<int32_t> wkn;
    // tensor-core (tcgen05) variant of the two GEMMs: operands kept as tf32 hi/lo pairs, K-major
    bool allow_tc = true;               // qp_set_option("tc", 0) keeps the fp32 FFMA GEMMs
    int tc_passes = 3;                  // 3 = 3xTF32 (fp32-grade), 1 = plain TF32
    bool tc_ready = false;
    DevBuf<float> p_hi[4], p_lo[4];     // price matrix shifted left by 0..3 elements (TMA 16-byte alignment)
    DevBuf<float> pT_hi, pT_lo, w1T_hi, w1T_lo, dzT_hi, dzT_lo, xe_hi, xe_lo;
    bool tc_a_lo_zero = false;          // the prices are tf32-exact: the A_lo tiles and MMAs are skipped
    int64_t tc_Kp = 0, tc_Bp = 0, tc_NsPad = 0;
    int tc_S = 1;
    CUtensorMap tm_p_hi[4], tm_p_lo[4], tm_pT_hi, tm_pT_lo, tm_w_hi, tm_w_lo, tm_dz_hi, tm_dz_lo, tm_xe_hi, tm_xe_lo;
    bool allow_fused = true;            // qp_set_option("fused", 0) forces the three-kernel path
    int rsag_mode = -1;                 // qp_set_option("rsag", 0 | 1): share exchange as all-to-all / reduce-scatter + all-gather; -1: by world size
    bool allow_nvls = true;             // qp_set_option("nvls", 0): unicast share exchange even where a multicast address exists
    DevBuf<long long> dbg;              // phase stamps of the persistent kernel (qp_debug_phase_clocks)
    bool want_dbg = false;
    bool want_fine = false;             // "phase_clocks" = 3: stamps inside the pair kernel's forward
    bool want_cta_ns = false;           // "phase_clocks" = 2: globaltimer stamps of every CTA instead (tensor-core kernel)
    bool coop_launch = false;
};

namespace {

int use_device(const qp_handle* h) {
    QP_CUDA(cudaSetDevice(h->device));
    return QP_OK;
}

size_t step_smem_bytes(const Hyper& hp, int mode, int nwarps) {
    const size_t rows_x = mode == kEnvStep ? hp.wn + 1 : hp.D;
    size_t f = (size_t)hp.D * hp.Hs + (size_t)hp.Hs * 4 + rows_x * kXStride;
    if (mode == kMsgUpdate) f += rows_x * kXStride;
    f += (size_t)nwarps * 2 * 3 * 32 + (size_t)hp.Hs * 3;
    f = (f + 3) & ~(size_t)3;
    return f * sizeof(float) + (size_t)((hp.D + 31) / 32) * sizeof(uint64_t);   // + one mbarrier per W1 row group
}

int make_geometry(const Hyper& hp, int max_smem, Geometry* g) {
    g->step_warps = hp.Hs / kColsPerWarp;
    for (int m = 0; m < 3; ++m) g->smem_step[m] = step_smem_bytes(hp, m, g->step_warps);
    const int H4 = hp.H / 4;
    g->n_rt = (hp.D + kDw1Rows - 1) / kDw1Rows;
    int T = std::max(1, (hp.H + kDw1TileCols - 1) / kDw1TileCols);
    while (T < H4 && (int64_t)g->n_rt * ((H4 + T - 1) / T) > kDw1MaxThreads) ++T;
    g->dw1_tiles = T;
    g->cg_max = (H4 + T - 1) / T;
    g->dw1_threads = (int)round_up((int64_t)g->n_rt * g->cg_max, 32);
    g->smem_dw1 = (size_t)kDw1Stages * kDw1ChunkEnvs *
                  ((size_t)g->n_rt * kDw1Rows + (size_t)g->cg_max * kDw1Cols) * sizeof(float);
    // large-batch K2
    g->big_n_rt = (hp.D + kDw1BigRows - 1) / kDw1BigRows;
    g->big_n_ct = (H4 + 1) / 2;
    g->big_threads = (int)round_up((int64_t)g->big_n_rt * g->big_n_ct, 32);
    g->smem_big = (size_t)kDw1Stages * kDw1ChunkEnvs *
                  ((size_t)g->big_n_rt * kDw1BigRows + (size_t)g->big_n_ct * kDw1BigCols) * sizeof(float);
    g->big_ok = g->big_threads <= kDw1MaxThreads && g->smem_big <= (size_t)max_smem;
    // persistent episode kernel
    g->ep_n_rt = (hp.wn + 1 + kEpRows - 1) / kEpRows;
    g->ep_n_ct = (H4 + 1) / 2;
    g->ep_threads = (int)std::max<int64_t>(32 * g->step_warps, round_up((int64_t)g->ep_n_rt * g->ep_n_ct, 32));
    {
        const size_t Dp = (size_t)g->ep_n_rt * kEpRows, Hp = (size_t)g->ep_n_ct * kEpCols;
        size_t f = (size_t)hp.D * hp.Hs + (size_t)hp.Hs * 4 + Dp * kXStride + 64 +
                   (size_t)g->step_warps * 2 * 3 * 32 + (size_t)hp.Hs * 3;
        f = (f + 3) & ~(size_t)3;
        g->smem_ep = f * sizeof(float) + (size_t)((hp.D + 31) / 32) * sizeof(uint64_t);
        g->ep_ok = g->ep_threads <= kEpMaxThreads && g->step_warps <= kMaxStepWarps &&
                   (int64_t)hp.wn * kTileEnvs <= 24LL * g->ep_threads &&
                   (size_t)kTileEnvs * Hp <= (size_t)hp.D * hp.Hs && g->smem_ep <= (size_t)max_smem;
    }
    if (g->step_warps > kMaxStepWarps || g->dw1_threads > kDw1MaxThreads ||
        g->smem_step[kMsgUpdate] > (size_t)max_smem || g->smem_dw1 > (size_t)max_smem)
        return fail(QP_EUNSUPPORTED,
                    "shape outside the fused fp32 path (needs W1 resident in shared memory: hidden <= 240, "
                    "(window+2)*hidden*4 + staging <= " + std::to_string(max_smem) + " bytes)");
    return QP_OK;
}

// large batches: one split per SM, the big-tile kernel
bool dw1_use_big(const Geometry& g, int64_t B, int num_sms) { return g.big_ok && B >= (int64_t)128 * num_sms; }

// episode splits of the dW1 grid: enough episodes per split to be worth a partial, all SMs used
int dw1_splits(int64_t B, int num_sms, int tiles) {
    return (int)std::max<int64_t>(1, std::min<int64_t>((B + 15) / 16, std::max(1, num_sms / tiles)));
}

// Host-side N(0,1) initialiser (Philox stream 2 + Box-Muller); stands in for the unseeded
// tf.RandomNormalInitializer() at QDecisionPolicyActor.scala:41, :45.
void gen_normal(uint64_t seed, int64_t n, float* out) {
    for (int64_t i = 0; i < n; ++i) {
        uint32_t r[4];
        philox4x32_10((uint32_t)i, (uint32_t)((uint64_t)i >> 32), 0u, 2u, (uint32_t)seed,
                      (uint32_t)(seed >> 32), r);
        const double u1 = ((double)(r[0] >> 8) + 0.5) * (1.0 / 16777216.0);
        const double u2 = ((double)(r[1] >> 8) + 0.5) * (1.0 / 16777216.0);
        out[i] = (float)(std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586476925286766559 * u2));
    }
}

int ensure_scratch(qp_handle* h, int64_t B, int* grid_step, int* grid_dw1) {
    const int64_t Bpad = round_up(B, kTileEnvs);
    const int64_t n_tiles = Bpad / kTileEnvs;
    const int g1 = (int)std::min<int64_t>(n_tiles, h->num_sms);
    const int g2 = dw1_use_big(h->geo, B, h->num_sms) ? h->num_sms : dw1_splits(B, h->num_sms, h->geo.dw1_tiles);
    QP_TRY(h->dz1.ensure((size_t)Bpad * h->hp.H));
    QP_TRY(h->xextra.ensure((size_t)Bpad * 2));
    QP_TRY(h->part1.ensure((size_t)g2 * h->P1));
    QP_TRY(h->part2.ensure((size_t)g1 * 3 * h->hp.H));
    h->Bpad = Bpad;
    *grid_step = g1;
    *grid_dw1 = g2;
    return QP_OK;
}

template <int MODE>
int launch_step(qp_handle* h, const StepArgs& a, int grid) {
    step_kernel<MODE><<<grid, h->geo.step_warps * 32, h->geo.smem_step[MODE], h->stream>>>(a);
    h->launches++;
    QP_CUDA(cudaGetLastError());
    return QP_OK;
}

template <int MODE>
int launch_dw1(qp_handle* h, const Dw1Args& a_in, int grid) {
    if (dw1_use_big(h->geo, a_in.B, h->num_sms)) {
        Dw1Args a = a_in;
        a.n_rt = h->geo.big_n_rt; a.cg_max = h->geo.big_n_ct;
        dw1_big_kernel<MODE><<<grid, h->geo.big_threads, h->geo.smem_big, h->stream>>>(a);
        h->launches++;
        QP_CUDA(cudaGetLastError());
        return QP_OK;
    }
    const Dw1Args& a = a_in;
    dw1_kernel<MODE><<<dim3(grid, h->geo.dw1_tiles), h->geo.dw1_threads, h->geo.smem_dw1, h->stream>>>(a);
    h->launches++;
    QP_CUDA(cudaGetLastError());
    return QP_OK;
}

int launch_apply(qp_handle* h, int n1, int n2, int do_reduce, int do_adagrad, int step_inc) {
    ApplyArgs a{};
    a.hp = h->hp;
    a.params = h->params.p; a.acc = h->acc.p; a.grad = h->grad.p;
    a.part1 = h->part1.p; a.n1 = n1; a.part2 = h->part2.p; a.n2 = n2;
    a.P = h->P; a.P1 = h->P1;
    a.do_reduce = do_reduce; a.do_adagrad = do_adagrad;
    a.d_step = h->d_step.p; a.step_inc = step_inc;
    const int64_t threads = h->P * kApplyLanes;
    const int blocks = (int)((threads + 255) / 256);
    apply_kernel<<<blocks, 256, 0, h->stream>>>(a);
    h->launches++;
    QP_CUDA(cudaGetLastError());
    return QP_OK;
}

// One env-step of every episode: K1, K2, K3.  ev != nullptr brackets each kernel with events.
int enqueue_env_step(qp_handle* h, const TraceDev* trace, int32_t trace_step0, bool timed) {
    int g1 = 0, g2 = 0;
    QP_TRY(ensure_scratch(h, h->n_envs, &g1, &g2));
    StepArgs s{};
    s.hp = h->hp;
    s.params = h->params.p;
    s.prices = h->prices.p; s.T = h->series_len; s.series_of_env = h->series_of_env.p;
    s.budget = h->budget.p; s.shares = h->shares.p; s.share_value = h->share_value.p;
    s.B0 = h->B0; s.N0 = h->N0; s.env_id_base = h->env_id_base; s.d_step = h->d_step.p;
    s.B = h->n_envs; s.Bpad = h->Bpad;
    s.dz1 = h->dz1.p; s.xextra = h->xextra.p; s.dw2_part = h->part2.p;
    if (trace) s.trace = *trace;
    s.trace_step0 = trace_step0;
    if (timed) QP_CUDA(cudaEventRecord(h->ev[0], h->stream));
    QP_TRY(launch_step<kEnvStep>(h, s, g1));
    if (timed) QP_CUDA(cudaEventRecord(h->ev[1], h->stream));

    Dw1Args d{};
    d.hp = h->hp;
    d.prices = h->prices.p; d.T = h->series_len; d.series_of_env = h->series_of_env.p; d.d_step = h->d_step.p;
    d.xextra = h->xextra.p; d.Bpad = h->Bpad; d.dz1 = h->dz1.p; d.B = h->n_envs; d.part = h->part1.p;
    d.n_rt = h->geo.n_rt; d.cg_max = h->geo.cg_max;
    QP_TRY(launch_dw1<kEnvStep>(h, d, g2));
    if (timed) QP_CUDA(cudaEventRecord(h->ev[2], h->stream));

    if (h->world > 1) {
        // sum the partials, all-reduce the P-float gradient over the ranks, then every rank applies
        // the identical sum: replicas stay bit-identical without any weight broadcast
        QP_TRY(launch_apply(h, g2, g1, 1, 0, 0));
        QP_NCCL(g_nccl.AllReduce(h->grad.p, h->grad.p, (size_t)h->P, kNcclFloat32, kNcclSum, h->comm, h->stream));
        QP_TRY(launch_apply(h, 0, 0, 0, 1, 1));
    } else {
        QP_TRY(launch_apply(h, g2, g1, 1, 1, 1));
    }
    if (timed) QP_CUDA(cudaEventRecord(h->ev[3], h->stream));
    return QP_OK;
}

// One env-step on the fan-out path: F1, F2, F3, AdaGrad (see qp_fanout.cuh).
int enqueue_env_step_fanout(qp_handle* h, const TraceDev* trace, int32_t trace_step0, bool timed = false) {
    const Hyper& hp = h->hp;
    QP_TRY(h->fan_zw.ensure((size_t)h->n_series * 2 * hp.H));
    QP_TRY(h->fan_part.ensure((size_t)h->fan_n_chunks * 6 * hp.H));
    QP_TRY(h->part1.ensure((size_t)kFanKSplit * h->P1));
    QP_TRY(h->part2.ensure((size_t)kFanKSplit * 3 * hp.H));
    FanArgs a{};
    a.hp = hp; a.params = h->params.p;
    a.prices = h->prices.p; a.T = h->series_len; a.n_series = h->n_series; a.d_step = h->d_step.p;
    a.budget = h->budget.p; a.shares = h->shares.p; a.share_value = h->share_value.p;
    a.B0 = h->B0; a.N0 = h->N0; a.env_id_base = h->env_id_base; a.B = h->n_envs;
    a.chunks = reinterpret_cast<const int4*>(h->fan_chunks.p); a.n_chunks = h->fan_n_chunks;
    a.zw = h->fan_zw.p; a.part = h->fan_part.p; a.part1 = h->part1.p; a.part2 = h->part2.p;
    if (trace) a.trace = *trace;
    a.trace_step0 = trace_step0;
    if (timed) QP_CUDA(cudaEventRecord(h->ev[0], h->stream));
    fan_zwin_kernel<<<dim3((unsigned)((h->n_series + kFanSeriesPerCta - 1) / kFanSeriesPerCta), (unsigned)((hp.H + 31) / 32)), 32 * kFanKSplit,
                      ((size_t)kFanSeriesPerCta * (hp.wn + 1) + (size_t)kFanKSplit * 2 * kFanSeriesPerCta * 32) * sizeof(float), h->stream>>>(a);
    QP_CUDA(cudaGetLastError());
    if (timed) QP_CUDA(cudaEventRecord(h->ev[1], h->stream));
    fan_row_kernel<<<h->fan_n_chunks, kFanChunk, ((size_t)2 * hp.H + kFanChunk) * sizeof(float4), h->stream>>>(a);
    QP_CUDA(cudaGetLastError());
    if (timed) QP_CUDA(cudaEventRecord(h->ev[2], h->stream));
    fan_grad_kernel<<<dim3((unsigned)((hp.H + kFanTile - 1) / kFanTile), (unsigned)((hp.wn + kFanTile - 1) / kFanTile + 1), kFanKSplit), 256, 0, h->stream>>>(a);
    QP_CUDA(cudaGetLastError());
    h->launches += 3;
    if (h->world > 1) {
        QP_TRY(launch_apply(h, kFanKSplit, kFanKSplit, 1, 0, 0));
        QP_NCCL(g_nccl.AllReduce(h->grad.p, h->grad.p, (size_t)h->P, kNcclFloat32, kNcclSum, h->comm, h->stream));
        QP_TRY(launch_apply(h, 0, 0, 0, 1, 1));
    } else {
        QP_TRY(launch_apply(h, kFanKSplit, kFanKSplit, 1, 1, 1));
    }
    if (timed) QP_CUDA(cudaEventRecord(h->ev[3], h->stream));
    return QP_OK;
}

// ---- tcgen05 GEMM plumbing ------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// The library links only the static runtime; the one driver call it needs is looked up at run time.
int get_encode_tiled(EncodeTiledFn* out) {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        QP_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
        if (!p || q != cudaDriverEntryPointSuccess) return fail(QP_ECUDA, "cuTensorMapEncodeTiled is not available in this driver");
        fn = (EncodeTiledFn)p;
    }
    *out = fn;
    return QP_OK;
}

// 2-D fp32 tensor [outer][inner] with row pitch `pitch_bytes`, boxes of 128 rows x 32 elements
// (128 B), 128-byte swizzle — the layout the K-major UMMA descriptor in qp_tc.cuh expects.
int make_tmap(CUtensorMap* m, const float* base, uint64_t inner, uint64_t outer, uint64_t pitch_bytes,
              uint32_t box_rows = kTcBM) {
    EncodeTiledFn enc = nullptr;
    QP_TRY(get_encode_tiled(&enc));
    const cuuint64_t dims[2] = {inner, outer};
    const cuuint64_t strides[1] = {pitch_bytes};
    const cuuint32_t box[2] = {kTcBK, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(QP_ECUDA, "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    return QP_OK;
}

// 3-D fp32 tensor [d2][d1][d0] for TMA stores of [32 x 32] boxes (128-byte rows, 128-byte swizzle)
int make_tmap3_store(CUtensorMap* m, const float* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t pitch1_bytes,
                     uint64_t pitch2_bytes) {
    EncodeTiledFn enc = nullptr;
    QP_TRY(get_encode_tiled(&enc));
    const cuuint64_t dims[3] = {d0, d1, d2};
    const cuuint64_t strides[2] = {pitch1_bytes, pitch2_bytes};
    const cuuint32_t box[3] = {32, 32, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)base, dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(QP_ECUDA, "cuTensorMapEncodeTiled (3-D) failed with code " + std::to_string((int)r));
    return QP_OK;
}

// 4-D fp32 tensor [d3][d2][d1][d0] for TMA stores of [32 (d2) x 32 (d0)] boxes (128-byte rows, 128-byte swizzle)
int make_tmap4_store(CUtensorMap* m, const float* base, const uint64_t dims_in[4], const uint64_t pitch_bytes[3], bool row_boxes = false) {
    EncodeTiledFn enc = nullptr;
    QP_TRY(get_encode_tiled(&enc));
    const cuuint64_t dims[4] = {dims_in[0], dims_in[1], dims_in[2], dims_in[3]};
    const cuuint64_t strides[3] = {pitch_bytes[0], pitch_bytes[1], pitch_bytes[2]};
    // row_boxes: [32 (d1) x 32 (d0)] boxes, rows of 128 bytes laid out linearly in shared memory (no swizzle)
    const cuuint32_t box[4] = {32, row_boxes ? 32u : 1u, row_boxes ? 1u : 32u, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (void*)base, dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, row_boxes ? CU_TENSOR_MAP_SWIZZLE_NONE : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(QP_ECUDA, "cuTensorMapEncodeTiled (4-D) failed with code " + std::to_string((int)r));
    return QP_OK;
}

int launch_tc_gemm(cudaStream_t st, const CUtensorMap& a_hi, const CUtensorMap& a_lo, const CUtensorMap& b_hi,
                   const CUtensorMap& b_lo, const TcGemmArgs& g, int tiles_m, int tiles_n, int splits) {
    static bool attr_set[64] = {false};
    int dev = 0;
    QP_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        QP_CUDA(cudaFuncSetAttribute(tc_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTcSmemBytes));
        attr_set[dev] = true;
    }
    tc_gemm_kernel<<<dim3(tiles_n, tiles_m, splits), 256, kTcSmemBytes, st>>>(a_hi, a_lo, b_hi, b_lo, g);
    QP_CUDA(cudaGetLastError());
    return QP_OK;
}

// QP_DEBUG_SYNC=1 in the environment synchronises after every launch of the GEMM path and names
// the kernel that failed (debug aid; never set for measurements).
int dbg_sync(qp_handle* h, const char* what) {
    static const bool on = std::getenv("QP_DEBUG_SYNC") != nullptr;
    if (!on) return QP_OK;
    cudaError_t e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) return fail(QP_ECUDA, std::string(what) + ": " + cudaGetErrorString(e));
    return QP_OK;
}

// One env-step on the GEMM-organised path: G1, E1, E2, G2, apply (see qp_wide.cuh).
int enqueue_env_step_wide(qp_handle* h, const TraceDev* trace, int32_t trace_step0, int64_t istep_host) {
    const Hyper& hp = h->hp;
    const int64_t B = h->n_envs, Bpad = round_up(B, kTileEnvs);
    const int tiles_m = (hp.D + kWgBM - 1) / kWgBM, tiles_n = (hp.H + kWgBN - 1) / kWgBN;
    const int S = (h->allow_tc && h->tc_ready)
                      ? h->tc_S
                      : (int)std::max<int64_t>(1, std::min<int64_t>({(int64_t)32, B / 256 > 0 ? B / 256 : 1,
                                                                     (h->num_sms * 6 + tiles_m * tiles_n - 1) / (tiles_m * tiles_n)}));
    const int colblocks = (hp.H + 255) / 256;
    const int S2 = (int)std::max<int64_t>(1, std::min<int64_t>(B / 64 > 0 ? B / 64 : 1, (h->num_sms * 2) / colblocks));
    QP_TRY(h->wZ1.ensure((size_t)2 * B * hp.H));
    QP_TRY(h->wdZ1.ensure((size_t)B * hp.H));
    QP_TRY(h->wdz2.ensure(B));
    QP_TRY(h->wkn.ensure(B));
    QP_TRY(h->xextra.ensure((size_t)Bpad * 2));
    QP_TRY(h->part1.ensure((size_t)S * h->P1));
    QP_TRY(h->part2.ensure((size_t)S2 * 3 * hp.H));
    WideArgs a{};
    a.hp = hp; a.params = h->params.p;
    a.prices = h->prices.p; a.T = h->series_len; a.series_of_env = h->series_of_env.p; a.d_step = h->d_step.p;
    a.budget = h->budget.p; a.shares = h->shares.p; a.share_value = h->share_value.p;
    a.B0 = h->B0; a.N0 = h->N0; a.env_id_base = h->env_id_base;
    a.B = B; a.Bpad = Bpad;
    a.Z1 = h->wZ1.p; a.dZ1 = h->wdZ1.p; a.xextra = h->xextra.p; a.kn = h->wkn.p; a.dz2 = h->wdz2.p;
    a.part1 = h->part1.p; a.part2 = h->part2.p; a.n_split = S; a.n_split2 = S2;
    if (trace) a.trace = *trace;
    a.trace_step0 = trace_step0;
    const bool tc = h->allow_tc && h->tc_ready;
    const int tiles_tc = hp.wn / kTcBM + ((hp.wn % kTcBM) > kWideMaxExtraRows - 2 ? 1 : 0);
    const int rows_tc = std::min(tiles_tc * kTcBM, hp.wn);
    if (tc) { a.xe_hi = h->xe_hi.p; a.xe_lo = h->xe_lo.p; a.xe_ld = h->tc_Bp; a.xe_row0 = rows_tc; }
    const int32_t istep = (int32_t)istep_host;
    if (tc) {
        // W1[0:wn, :] -> K-major tf32 hi/lo pair [H, Kp] (columns wn..Kp-1 stay zero)
        tc_transpose_split_kernel<<<dim3((hp.H + 31) / 32, (hp.wn + 31) / 32), dim3(32, 8), 0, h->stream>>>(
            h->params.p, hp.wn, hp.H, hp.H, h->w1T_hi.p, h->w1T_lo.p, h->tc_Kp);
        QP_CUDA(cudaGetLastError());
        QP_TRY(dbg_sync(h, "W1 transpose/split"));
        for (int half = 0; half < 2; ++half) {      // state s, then s' (window shifted by one)
            TcGemmArgs g{};
            const int shift = (istep + half) & 3;       // copy whose window start is 16-byte aligned
            g.a_k0 = istep + half - shift; g.a_row0 = 0; g.b_k0 = 0; g.b_row0 = 0;
            g.nkb = (int)(h->tc_Kp / kTcBK); g.passes = h->tc_passes; g.a_lo_zero = h->tc_a_lo_zero ? 1 : 0;
            g.C = h->wZ1.p + (size_t)half * B * hp.H; g.ldc = hp.H; g.c_split_stride = 0;
            g.m_valid = (int)B; g.n_valid = hp.H;
            QP_TRY(launch_tc_gemm(h->stream, h->tm_p_hi[shift], h->tm_p_lo[shift], h->tm_w_hi, h->tm_w_lo, g,
                                  (int)((B + kTcBM - 1) / kTcBM), (hp.H + kTcBN - 1) / kTcBN, 1));
            QP_TRY(dbg_sync(h, half ? "tc G1 (s')" : "tc G1 (s)"));
        }
        h->launches += 3;
    } else {
        wide_gemm_kernel<0><<<dim3((unsigned)((2 * B + kWgBM - 1) / kWgBM), tiles_n), kWgThreads, 0, h->stream>>>(a);
        QP_CUDA(cudaGetLastError());
        h->launches += 1;
    }
    const int row_blocks = (int)std::min<int64_t>((B + 7) / 8, (int64_t)h->num_sms * 4);
    wide_row_kernel<0><<<row_blocks, 256, (size_t)hp.H * 6 * sizeof(float), h->stream>>>(a);
    QP_CUDA(cudaGetLastError());
    wide_dw2_kernel<<<dim3(colblocks, S2), 256, 0, h->stream>>>(a);
    QP_CUDA(cudaGetLastError());
    QP_TRY(dbg_sync(h, "row / dW2 kernels"));
    if (tc) {
        // dZ1 [B, H] -> K-major (episodes contiguous) tf32 hi/lo pair [H, Bp]
        tc_transpose_split_kernel<<<dim3((hp.H + 31) / 32, (unsigned)((B + 31) / 32)), dim3(32, 8), 0, h->stream>>>(
            h->wdZ1.p, B, hp.H, hp.H, h->dzT_hi.p, h->dzT_lo.p, h->tc_Bp);
        QP_CUDA(cudaGetLastError());
        QP_TRY(dbg_sync(h, "dZ1 transpose/split"));
        const int64_t eps = h->tc_Bp / S;
        if (tiles_tc > 0) {
            TcGemmArgs g{};
            g.a_k0 = 0; g.a_row0 = istep; g.b_k0 = 0; g.b_row0 = 0;
            g.nkb = (int)(eps / kTcBK); g.passes = h->tc_passes; g.a_lo_zero = h->tc_a_lo_zero ? 1 : 0;
            g.C = h->part1.p; g.ldc = hp.H; g.c_split_stride = h->P1; g.m_valid = rows_tc; g.n_valid = hp.H;
            QP_TRY(launch_tc_gemm(h->stream, h->tm_pT_hi, h->tm_pT_lo, h->tm_dz_hi, h->tm_dz_lo, g, tiles_tc,
                                  (hp.H + kTcBN - 1) / kTcBN, S));
            QP_TRY(dbg_sync(h, "tc G2"));
        }
        {   // the remaining rows of dW1 (window rows past the last full tile, budget, shares): one more tile
            TcGemmArgs g{};
            g.a_k0 = 0; g.a_row0 = 0; g.b_k0 = 0; g.b_row0 = 0;
            g.nkb = (int)(eps / kTcBK); g.passes = h->tc_passes; g.a_lo_zero = 0;
            g.C = h->part1.p + (size_t)rows_tc * hp.H; g.ldc = hp.H; g.c_split_stride = h->P1;
            g.m_valid = hp.D - rows_tc; g.n_valid = hp.H;
            QP_TRY(launch_tc_gemm(h->stream, h->tm_xe_hi, h->tm_xe_lo, h->tm_dz_hi, h->tm_dz_lo, g, 1,
                                  (hp.H + kTcBN - 1) / kTcBN, S));
            QP_TRY(dbg_sync(h, "tc G2 (extra rows)"));
        }
        h->launches += 5;
    } else {
        wide_gemm_kernel<1><<<dim3(tiles_m, tiles_n, S), kWgThreads, 0, h->stream>>>(a);
        QP_CUDA(cudaGetLastError());
        h->launches += 3;
    }
    if (h->world > 1) {
        QP_TRY(launch_apply(h, S, S2, 1, 0, 0));
        QP_NCCL(g_nccl.AllReduce(h->grad.p, h->grad.p, (size_t)h->P, kNcclFloat32, kNcclSum, h->comm, h->stream));
        QP_TRY(launch_apply(h, 0, 0, 0, 1, 1));
    } else {
        QP_TRY(launch_apply(h, S, S2, 1, 1, 1));
    }
    return QP_OK;
}

// Builds what the tcgen05 GEMMs need for the current episodes: tf32 hi/lo pairs of the price matrix
// (row-major for G1, transposed for G2), zeroed K-major buffers for W1^T and dZ1^T, and the TMA maps.
int wide_prepare_tc(qp_handle* h) {
    const Hyper& hp = h->hp;
    const int64_t Ns = h->n_series, T = h->series_len, B = h->n_envs;
    const int64_t NsPad = round_up(Ns, 4);
    h->tc_NsPad = NsPad;
    h->tc_Kp = round_up(hp.wn, kTcBK);
    // episode splits of the backward GEMM: fill the SMs with whole waves
    const int tiles = std::max(1, hp.wn / kTcBM + ((hp.wn % kTcBM) > kWideMaxExtraRows - 2 ? 1 : 0)) * ((hp.H + kTcBN - 1) / kTcBN);
    int best = 1; double best_eff = 0.0;
    for (int S = 1; S <= 32; ++S) {
        if (B / S < 256 && S > 1) break;
        const int ctas = tiles * S;
        const double eff = (double)ctas / ((double)((ctas + h->num_sms - 1) / h->num_sms) * h->num_sms);
        if (eff > best_eff + 1e-9) { best_eff = eff; best = S; }
    }
    h->tc_S = best;
    h->tc_Bp = round_up(B, (int64_t)kTcBK * best);
    for (int j = 0; j < 4; ++j) { QP_TRY(h->p_hi[j].ensure((size_t)Ns * T)); QP_TRY(h->p_lo[j].ensure((size_t)Ns * T)); }
    QP_TRY(h->pT_hi.ensure((size_t)T * NsPad)); QP_TRY(h->pT_lo.ensure((size_t)T * NsPad));
    QP_TRY(h->w1T_hi.ensure((size_t)hp.H * h->tc_Kp)); QP_TRY(h->w1T_lo.ensure((size_t)hp.H * h->tc_Kp));
    QP_TRY(h->dzT_hi.ensure((size_t)hp.H * h->tc_Bp)); QP_TRY(h->dzT_lo.ensure((size_t)hp.H * h->tc_Bp));
    QP_TRY(h->xe_hi.ensure((size_t)kWideMaxExtraRows * h->tc_Bp)); QP_TRY(h->xe_lo.ensure((size_t)kWideMaxExtraRows * h->tc_Bp));
    QP_CUDA(cudaMemsetAsync(h->xe_hi.p, 0, (size_t)kWideMaxExtraRows * h->tc_Bp * sizeof(float), h->stream));
    QP_CUDA(cudaMemsetAsync(h->xe_lo.p, 0, (size_t)kWideMaxExtraRows * h->tc_Bp * sizeof(float), h->stream));
    QP_CUDA(cudaMemsetAsync(h->pT_hi.p, 0, (size_t)T * NsPad * sizeof(float), h->stream));
    QP_CUDA(cudaMemsetAsync(h->pT_lo.p, 0, (size_t)T * NsPad * sizeof(float), h->stream));
    QP_CUDA(cudaMemsetAsync(h->w1T_hi.p, 0, (size_t)hp.H * h->tc_Kp * sizeof(float), h->stream));
    QP_CUDA(cudaMemsetAsync(h->w1T_lo.p, 0, (size_t)hp.H * h->tc_Kp * sizeof(float), h->stream));
    QP_CUDA(cudaMemsetAsync(h->dzT_hi.p, 0, (size_t)hp.H * h->tc_Bp * sizeof(float), h->stream));
    QP_CUDA(cudaMemsetAsync(h->dzT_lo.p, 0, (size_t)hp.H * h->tc_Bp * sizeof(float), h->stream));
    for (int j = 0; j < 4; ++j)
        tc_split_kernel<<<(unsigned)((Ns * T + 255) / 256), 256, 0, h->stream>>>(h->prices.p, h->p_hi[j].p, h->p_lo[j].p, Ns * T, T, j);
    tc_transpose_split_kernel<<<dim3((unsigned)((T + 31) / 32), (unsigned)((Ns + 31) / 32)), dim3(32, 8), 0, h->stream>>>(
        h->prices.p, Ns, T, T, h->pT_hi.p, h->pT_lo.p, NsPad);
    QP_CUDA(cudaGetLastError());
    h->launches += 5;
    for (int j = 0; j < 4; ++j) {
        QP_TRY(make_tmap(&h->tm_p_hi[j], h->p_hi[j].p, T, Ns, T * sizeof(float)));
        QP_TRY(make_tmap(&h->tm_p_lo[j], h->p_lo[j].p, T, Ns, T * sizeof(float)));
    }
    QP_TRY(make_tmap(&h->tm_pT_hi, h->pT_hi.p, NsPad, T, NsPad * sizeof(float)));
    QP_TRY(make_tmap(&h->tm_pT_lo, h->pT_lo.p, NsPad, T, NsPad * sizeof(float)));
    QP_TRY(make_tmap(&h->tm_w_hi, h->w1T_hi.p, h->tc_Kp, hp.H, h->tc_Kp * sizeof(float)));
    QP_TRY(make_tmap(&h->tm_w_lo, h->w1T_lo.p, h->tc_Kp, hp.H, h->tc_Kp * sizeof(float)));
    QP_TRY(make_tmap(&h->tm_dz_hi, h->dzT_hi.p, h->tc_Bp, hp.H, h->tc_Bp * sizeof(float)));
    QP_TRY(make_tmap(&h->tm_dz_lo, h->dzT_lo.p, h->tc_Bp, hp.H, h->tc_Bp * sizeof(float)));
    QP_TRY(make_tmap(&h->tm_xe_hi, h->xe_hi.p, h->tc_Bp, kWideMaxExtraRows, h->tc_Bp * sizeof(float)));
    QP_TRY(make_tmap(&h->tm_xe_lo, h->xe_lo.p, h->tc_Bp, kWideMaxExtraRows, h->tc_Bp * sizeof(float)));
    {   // are the prices tf32-exact (e.g. quarter ticks below 2048)?  Then A_lo == 0 exactly and is skipped.
        DevBuf<int32_t> tmp;
        QP_TRY(tmp.ensure(1));
        int* flag = tmp.p;
        QP_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), h->stream));
        tc_any_nonzero_kernel<<<(unsigned)((Ns * T + 255) / 256), 256, 0, h->stream>>>(h->p_lo[0].p, Ns * T, flag);
        int host_flag = 1;
        QP_CUDA(cudaMemcpyAsync(&host_flag, flag, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        QP_CUDA(cudaStreamSynchronize(h->stream));
        tmp.release();
        h->tc_a_lo_zero = host_flag == 0;
        h->launches += 1;
    }
    h->tc_ready = true;
    return QP_OK;
}

size_t p2p_tc_offset(const qp_handle* h) { return (size_t)round_up(256 + (int64_t)2 * h->P * (int64_t)sizeof(float), 256); }
// (the pair kernel indexes its shares in the transposed space [column][row padded to 4] + W2)
size_t p2p_pad(const qp_handle* h) { return (size_t)round_up(std::max<int64_t>(h->P, (int64_t)h->hp.H * c2_kk(h->hp.D) + 3 * h->hp.H), 64); }

// what both tensor-core persistent kernels need: the shape, the buffers, and (several GPUs) the flow form of phase B
bool ep_tc_base(const qp_handle* h) { return h->allow_ep_tc && h->ep_tc_ready && (h->world == 1 || h->et_flow_ok); }
// the one-tile-per-CTA kernel feeds the state matrix to the tensor core as it is: tf32-exact prices only
bool ep_tc_usable(const qp_handle* h) {
    const bool local = ep_tc_base(h) && h->prices_exact;
    return h->world == 1 ? local : (local && h->agree_valid && h->agree_ep_tc);
}

// the pair kernel: one 64-episode tile per CTA pair (no waves), everything episode_tc_kernel needs plus its own shape limits
// (prices that are not tf32-exact: 32 real + 32 virtual episodes per pair, see episode_c2_kernel's kSplit)
int64_t c2_envs_per_pair(const qp_handle* h) { return h->prices_exact ? kC2Envs : kC2Envs / 2; }
bool c2_one_tile(const qp_handle* h) { return h->n_envs <= c2_envs_per_pair(h) * (h->c2_grid / 2); }
// more tiles than pairs: the pair kernel in waves (tf32-exact prices only; option "ep_c2_waves")
bool ep_c2_local(const qp_handle* h) {
    return ep_tc_base(h) && h->allow_ep_c2 && h->ep_c2_ready && (c2_one_tile(h) || (h->prices_exact && h->allow_c2_waves));
}
bool ep_c2_usable(const qp_handle* h) { return ep_c2_local(h) && (h->world == 1 || (h->agree_valid && h->agree_c2)); }
// several GPUs: the share-exchange scheme and the multicast stores, as ALL ranks have them (agreed at qp_env_create)
bool c2_rsag_local(const qp_handle* h) { return h->rsag_mode == 1 || (h->rsag_mode < 0 && h->world >= 8); }
bool c2_nvls_local(const qp_handle* h) { return h->p2p_mc != nullptr && h->allow_nvls; }
bool c2_rsag(const qp_handle* h) { return h->world > 1 && c2_rsag_local(h) && h->agree_valid && h->agree_rsag; }
bool c2_nvls(const qp_handle* h) { return h->world > 1 && c2_nvls_local(h) && h->agree_valid && h->agree_nvls; }

// Which kernels serve qp_env_step.  With several GPUs every input of the decision is a fact ALL ranks share
// (agreed at qp_env_create): the three multi-GPU step protocols do not interoperate.
bool use_fused(const qp_handle* h) {
    const bool single = h->world == 1;
    const bool peers = h->world > 1 && h->world <= 8 && h->p2p_ready && h->allow_p2p && h->agree_valid && h->agree_p2p;
    // one tile per SM for the FFMA kernel; the tensor-core kernel runs any number of tiles in waves
    const bool fits_local = h->n_envs <= (int64_t)kTileEnvs * h->num_sms;
    const bool fits = single ? fits_local : (fits_local && h->agree_fits);
    const bool fused_ok = !h->wide && h->allow_fused && h->coop_launch && h->geo.ep_ok && (single || h->agree_fused);
    // (the FFMA form of the persistent kernel, 52-54 us per env-step at the reference shape against 72 for the three-kernel
    // path, serves what fits one tile per SM and no tensor-core kernel takes)
    return fused_ok && (single || peers) && (fits || ep_tc_usable(h) || ep_c2_usable(h));
}

// The fan-out path (qp_fanout.cuh): episodes grouped by series.  Taken on its own where it pays: at least eight episodes
// per series on average and no pair kernel for the batch (the one-tile-per-CTA kernel in waves and the three-kernel path
// both redo the window contraction per episode).  Several GPUs: only when asked for ("fanout" = 2 on every rank), with
// the NCCL all-reduce between the gradient sum and AdaGrad.
bool use_fanout(const qp_handle* h) {
    if (h->wide || !h->has_env || !h->fan_ok || h->allow_fanout == 0 || h->hp.H > kFanChunk) return false;
    if (h->allow_fanout >= 2) return true;
    return h->world == 1 && h->n_envs >= 8 * h->n_series && !(use_fused(h) && ep_c2_usable(h) && c2_one_tile(h));
}

// dst[c][r] = src[r][c] (dst rows of ld_dst floats, the columns r >= R left as they are: zero-filled by the caller)
__global__ void transpose_kernel(const float* __restrict__ src, int64_t R, int64_t C, float* __restrict__ dst, int64_t ld_dst) {
    __shared__ float tile[32][33];
    const int64_t r0 = (int64_t)blockIdx.y * 32, c0 = (int64_t)blockIdx.x * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int64_t r = r0 + i, c = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (r < R && c < C) ? src[r * C + c] : 0.0f;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int64_t c = c0 + i, r = r0 + threadIdx.x;
        if (c < C && r < R) dst[c * ld_dst + r] = tile[threadIdx.x][i];
    }
}

// The pair kernel in waves fetches a tile's window with TMA where episode e is series e: needs the transposed price matrix.
int c2_window_prepare(qp_handle* h) {
    if (h->c2_win_ready) return QP_OK;
    const int64_t T = h->series_len, Ns = h->n_series, ld = round_up(Ns, 64);
    QP_TRY(h->c2_pT.ensure((size_t)T * ld));
    QP_CUDA(cudaMemsetAsync(h->c2_pT.p, 0, (size_t)T * ld * sizeof(float), h->stream));
    transpose_kernel<<<dim3((unsigned)((T + 31) / 32), (unsigned)((Ns + 31) / 32)), dim3(32, 8), 0, h->stream>>>(h->prices.p, Ns, T, h->c2_pT.p, ld);
    QP_CUDA(cudaGetLastError());
    h->launches += 1;
    QP_TRY(make_tmap(&h->tm_c2_win, h->c2_pT.p, (uint64_t)ld, (uint64_t)T, (uint64_t)ld * sizeof(float), (uint32_t)(h->hp.wn + 1)));
    h->c2_win_ready = true;
    return QP_OK;
}

// The pair kernel (qp_ep_c2.cuh) on top of what ep_tc_prepare has set up; leaves ep_c2_ready false when the shape or
// the device does not take it.
int ep_c2_prepare(qp_handle* h) {
    h->ep_c2_ready = false;
    const Hyper& hp = h->hp;
    if (hp.H % 8 != 0 || !h->coop_launch) return QP_OK;
    const int Hh = hp.H / 2, Nh = (int)round_up(Hh, 8), grid = h->num_sms & ~1;
    const C2Smem sm = c2_smem(Hh, Nh);
    // the window buffer holds kC2Slide rows of slack behind window element wn + 1; the copy-back moves (wn + 1) * 16
    // float4 with 12 per thread; the epilogue stages 8 KB per warp over regA + regW; Nh <= 128 TMEM columns per accumulator
    if (hp.wn + 2 + kC2Slide > kC2Rows || (hp.wn + 1) * 16 > 12 * kC2Threads || Nh > 128 || 2 * Hh > kC2Threads ||
        sm.total + 1024 > (size_t)h->max_smem || grid < 8)
        return QP_OK;
    {   // phase B's bulk-copy branch must apply (the kernel has no other)
        const int n_w1 = std::max(1, grid - 2);
        const int64_t P1T = (int64_t)hp.H * c2_kk(hp.D);       // the pair kernel's partial is [column][row padded to 4]
        const int chunk = (int)(((P1T + n_w1 - 1) / n_w1 + 3) & ~(int64_t)3);
        const int chunk2 = ((3 * hp.H + 1) / 2 + 3) & ~3;
        const int pitch = std::max(chunk, chunk2);
        // the sum: (sub, float4 group) threads, at most 16 partials per thread; the sub sums meet in [8][pitch] floats
        const int ng4 = pitch / 4, nsub = std::min(8, kC2Threads / std::max(1, ng4));
        if (pitch > kC2Threads || nsub < 1 || (grid / 2 + nsub - 1) / nsub > 16 || (size_t)8 * pitch * 4 > sm.op_bytes) return QP_OK;
    }
    for (auto* k : {episode_c2_kernel<0, false, false>, episode_c2_kernel<1, false, false>, episode_c2_kernel<2, false, false>,
                    episode_c2_kernel<0, true, false>, episode_c2_kernel<1, true, false>, episode_c2_kernel<2, true, false>,
                    episode_c2_kernel<0, false, true>, episode_c2_kernel<1, false, true>, episode_c2_kernel<2, false, true>})
        QP_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sm.total + 1024)));
    {   // all pairs must be co-resident (cooperative launch of clusters of two)
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kC2Threads); cfg.dynamicSmemBytes = sm.total + 1024; cfg.stream = h->stream;
        int n_clusters = 0;
        if (cudaOccupancyMaxActiveClusters(&n_clusters, episode_c2_kernel<1, false, true>, &cfg) != cudaSuccess) { cudaGetLastError(); return QP_OK; }
        if (n_clusters * 2 < grid) return QP_OK;
    }
    QP_TRY(make_tmap(&h->tm_c2_hi, h->et_w_hi.p, h->et_Kp, hp.H, (uint64_t)h->et_Kp * sizeof(float), (uint32_t)Nh));
    QP_TRY(make_tmap(&h->tm_c2_lo, h->et_w_lo.p, h->et_Kp, hp.H, (uint64_t)h->et_Kp * sizeof(float), (uint32_t)Nh));
    h->c2_Hh = Hh; h->c2_Nh = Nh; h->c2_grid = grid; h->c2_smem_bytes = sm.total + 1024;
    h->ep_c2_ready = true;
    return QP_OK;
}

// Everything the tensor-core persistent kernel needs for the current episodes; leaves ep_tc_ready
// false (and the FFMA persistent kernel in charge) when a precondition does not hold.
int ep_tc_prepare(qp_handle* h, bool* flag_pending) {
    h->ep_tc_ready = false;
    h->et_flow_ok = false;
    *flag_pending = false;
    const Hyper& hp = h->hp;
    if (!h->tile_ok || !h->geo.ep_ok || hp.H % 4 != 0 || hp.H > kEtThreads) return QP_OK;
    const int Kp = (int)round_up(hp.wn, 32), nkb = Kp / 32, Np = (int)round_up(hp.H, 16), halves = (hp.wn + 127) / 128;
    const size_t TB = (size_t)Np * 128;
    const size_t smem = 1024 + (size_t)nkb * 8192 + 4 * TB + (size_t)halves * 16384 + 16 + (size_t)hp.H * 16 +
                        (size_t)hp.H * 8 + 256 + (size_t)kEtWarps * 384 * 4 + 256 + 8 + 256;
    // the window ring needs one free row beyond the window; the partial's boxes are staged in the operand ring
    // (8 KB per warp) and, for the warps that do not fit there, behind h(s) in the A-operand region (4 KB each)
    const size_t n_ring = std::min<size_t>(kEtWarps, 4 * TB / 8192);
    const size_t stage_end = round_up((int64_t)Np * kXStride * 4, 1024) + (kEtWarps - n_ring) * 4096;
    if (halves * 128 < hp.wn + 2 || stage_end > (size_t)nkb * 8192 || ((size_t)hp.H * 4) % 16 != 0 || ((size_t)h->P1 * 4) % 16 != 0)
        return QP_OK;
    if (Np > 256 || halves > 2 || halves * Np > 512 || (size_t)nkb * 8192 < 2 * (size_t)Np * kXStride * 4 || 
        smem > (size_t)h->max_smem)
        return QP_OK;
    // the state matrix goes to the tensor core as its tf32 value alone: the prices must be tf32-exact.  The flag lives
    // in a handle-owned buffer (no allocation here) and its read-back rides on the synchronisation qp_env_reset makes
    // anyway: see ep_tc_finish.
    QP_CUDA(cudaMemsetAsync(h->flags4.p, 0, 4 * sizeof(int32_t), h->stream));
    const int64_t n = h->n_series * h->series_len;
    tc_inexact_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(h->prices.p, n, h->flags4.p);
    h->launches += 1;
    *h->host_flag = 1;
    QP_CUDA(cudaMemcpyAsync(h->host_flag, h->flags4.p, sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    *flag_pending = true;
    {   // phase B's bulk-copy branch ("flow") is the one that carries the NVLink exchange: same arithmetic as the kernel
        const int64_t P1 = h->P1;
        const int n_w1 = std::max(1, h->num_sms - 2);
        const int chunk = (int)(((P1 + n_w1 - 1) / n_w1 + 3) & ~(int64_t)3);
        const int chunk2 = ((3 * hp.H + 1) / 2 + 3) & ~3;
        const int pitch = std::max(chunk, chunk2);
        const size_t off_xs = (size_t)nkb * 8192 + 4 * TB;
        const int rows_cap = std::min(128, (int)(off_xs / ((size_t)pitch * 4)) & ~3);
        h->et_flow_ok = h->num_sms >= 8 && pitch <= kEtThreads && rows_cap >= 4;
    }
    QP_TRY(h->et_w_hi.ensure((size_t)hp.H * Kp)); QP_TRY(h->et_w_lo.ensure((size_t)hp.H * Kp));
    QP_CUDA(cudaMemsetAsync(h->et_w_hi.p, 0, (size_t)hp.H * Kp * sizeof(float), h->stream));
    QP_CUDA(cudaMemsetAsync(h->et_w_lo.p, 0, (size_t)hp.H * Kp * sizeof(float), h->stream));
    QP_TRY(make_tmap(&h->tm_et_hi, h->et_w_hi.p, Kp, hp.H, (uint64_t)Kp * sizeof(float), (uint32_t)Np));
    QP_TRY(make_tmap(&h->tm_et_lo, h->et_w_lo.p, Kp, hp.H, (uint64_t)Kp * sizeof(float), (uint32_t)Np));
    h->et_Kp = Kp; h->et_nkb = nkb; h->et_Np = Np; h->et_halves = halves; h->et_smem = smem;
    h->ep_tc_ready = true;      // provisional: ep_tc_finish takes it back if a price is not tf32-exact
    QP_TRY(ep_c2_prepare(h));
    return QP_OK;
}

// Second half of the preparation, called once the stream has been synchronised (qp_env_reset does that for
// everything qp_env_create enqueued): read the exactness flag, and (several GPUs) agree on the execution path
// with the other ranks.
int ep_tc_finish(qp_handle* h, bool flag_pending) {
    h->prices_exact = !(flag_pending && *h->host_flag != 0);
    h->agree_valid = false;
    if (h->world > 1 && h->comm) {
        const int32_t mine[8] = {
            (h->allow_ep_tc && h->ep_tc_ready && h->et_flow_ok && h->prices_exact) ? 1 : 0,
            (h->n_envs <= (int64_t)kTileEnvs * h->num_sms) ? 1 : 0,
            (!h->wide && h->allow_fused && h->coop_launch && h->geo.ep_ok) ? 1 : 0,
            (h->p2p_ready && h->allow_p2p) ? 1 : 0,
            ep_c2_local(h) ? 1 : 0, c2_rsag_local(h) ? 1 : 0, c2_nvls_local(h) ? 1 : 0, 0};
        int32_t all[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        QP_CUDA(cudaMemcpyAsync(h->flags4.p, mine, sizeof mine, cudaMemcpyHostToDevice, h->stream));
        QP_NCCL(g_nccl.AllReduce(h->flags4.p, h->flags4.p, 8, kNcclInt32, kNcclMin, h->comm, h->stream));
        QP_CUDA(cudaMemcpyAsync(all, h->flags4.p, sizeof all, cudaMemcpyDeviceToHost, h->stream));
        QP_CUDA(cudaStreamSynchronize(h->stream));
        h->agree_ep_tc = all[0] != 0; h->agree_fits = all[1] != 0; h->agree_fused = all[2] != 0; h->agree_p2p = all[3] != 0;
        h->agree_c2 = all[4] != 0; h->agree_rsag = all[5] != 0; h->agree_nvls = all[6] != 0;
        h->agree_valid = true;
    }
    return QP_OK;
}

// n_steps env-steps of every episode in ONE cooperative launch of the persistent episode kernel.
int enqueue_env_steps_fused(qp_handle* h, int32_t n_steps, const TraceDev* trace, int32_t trace_step0,
                            bool apply_in_kernel = true, int32_t step_offset = 0) {
    if (n_steps <= 0) return QP_OK;
    const int64_t Bpad = round_up(h->n_envs, kTileEnvs);
    const int64_t n_tiles = Bpad / kTileEnvs;
    QP_TRY(h->part1.ensure((size_t)n_tiles * (size_t)((int64_t)h->hp.H * c2_kk(h->hp.D))));      // (the pair kernel's partials: [column][row padded to 4])
    QP_TRY(h->part2.ensure((size_t)n_tiles * 3 * h->hp.H));
    EpisodeArgs ea{};
    StepArgs& s = ea.s;
    s.hp = h->hp;
    s.params = h->params.p;
    s.prices = h->prices.p; s.T = h->series_len; s.series_of_env = h->series_of_env.p;
    s.budget = h->budget.p; s.shares = h->shares.p; s.share_value = h->share_value.p;
    s.B0 = h->B0; s.N0 = h->N0; s.env_id_base = h->env_id_base; s.d_step = h->d_step.p;
    s.B = h->n_envs; s.Bpad = Bpad; s.dw2_part = h->part2.p;
    if (trace) s.trace = *trace;
    s.trace_step0 = trace_step0;
    ea.step0 = (int32_t)h->steps_done + step_offset; ea.n_steps = n_steps;
    ea.apply_in_kernel = apply_in_kernel ? 1 : 0;
    ea.world = 1; ea.rank = 0;
    if (apply_in_kernel && h->world > 1) {     // in-kernel all-reduce over peer memory
        ea.world = h->world; ea.rank = h->rank; ea.flag_base = h->p2p_flag_base;
        for (int r = 0; r < h->world; ++r) {
            ea.flags[r] = reinterpret_cast<unsigned int*>(h->p2p_peer_block[r]);
            ea.xg[r] = reinterpret_cast<float*>(reinterpret_cast<char*>(h->p2p_peer_block[r]) + 256);
        }
        h->p2p_flag_base += (unsigned int)n_steps;
    }
    ea.acc = h->acc.p; ea.grad = h->grad.p; ea.params_rw = h->params.p; ea.part1 = h->part1.p;
    ea.d_step_rw = h->d_step.p;
    ea.n_rt = h->geo.ep_n_rt; ea.n_ct = h->geo.ep_n_ct; ea.nwf = h->geo.step_warps;
    ea.dbg = nullptr;
    if (h->want_dbg) {
        QP_TRY(h->dbg.ensure((size_t)n_steps * qp::kDbgSlots));
        ea.dbg = h->dbg.p;
    }
    if (ep_c2_usable(h) && apply_in_kernel) {
        // CTA pairs that split the hidden dimension: one 64-episode tile per pair
        const int64_t n_pairs = (h->n_envs + c2_envs_per_pair(h) - 1) / c2_envs_per_pair(h);
        EpC2Args x{};
        x.split = h->prices_exact ? 0 : 1;
        x.ea = ea;
        x.w1T_hi = h->et_w_hi.p; x.w1T_lo = h->et_w_lo.p;
        x.Kp = h->et_Kp; x.nkb = h->et_nkb; x.Hh = h->c2_Hh; x.Nh = h->c2_Nh; x.dbg_mode = h->c2_dbg;
        QP_TRY(h->et_published.ensure(2));
        x.published = h->et_published.p;
        x.p_pad = (int64_t)p2p_pad(h);
        for (int r = 0; r < ea.world && ea.world > 1; ++r)
            x.pdata[r] = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(h->p2p_peer_block[r]) + p2p_tc_offset(h));
        x.rsag = (ea.world > 1 && c2_rsag(h)) ? 1 : 0;
        x.mc_pdata = (ea.world > 1 && c2_nvls(h))
                         ? reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(h->p2p_mc) + p2p_tc_offset(h)) : nullptr;
        if (h->want_cta_ns) {        // [n_steps][grid][8]
            QP_TRY(h->dbg.ensure((size_t)n_steps * h->c2_grid * 8));
            QP_CUDA(cudaMemsetAsync(h->dbg.p, 0, (size_t)n_steps * h->c2_grid * 8 * sizeof(long long), h->stream));
            x.ea.dbg = nullptr;
            x.cta_ns = h->dbg.p;
        }
        if (h->want_fine) {          // fine stamps share the dbg buffer: [n_steps][64]
            QP_TRY(h->dbg.ensure((size_t)n_steps * 64));
            QP_CUDA(cudaMemsetAsync(h->dbg.p, 0, (size_t)n_steps * 64 * sizeof(long long), h->stream));
            x.ea.dbg = nullptr;
            x.fine = h->dbg.p;
        }
        const int64_t n_parts = std::min<int64_t>(n_pairs, h->c2_grid / 2);          // one partial per pair
        if (h->tm_c2_part_ptr != h->part1.p || h->tm_c2_part_pairs != n_parts) {
            // [pair][half][column][row]: rows end at wn, columns at Hh (the boxes are clipped there)
            const uint64_t Kk = (uint64_t)c2_kk(h->hp.D);
            const uint64_t dims[4] = {(uint64_t)h->hp.wn, (uint64_t)h->c2_Hh, 2, (uint64_t)n_parts};
            const uint64_t pitch[3] = {Kk * sizeof(float), (uint64_t)h->c2_Hh * Kk * sizeof(float), (uint64_t)h->hp.H * Kk * sizeof(float)};
            QP_TRY(make_tmap4_store(&h->tm_c2_part, h->part1.p, dims, pitch, /*row_boxes=*/true));
            h->tm_c2_part_ptr = h->part1.p; h->tm_c2_part_pairs = n_parts;
        }
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(h->c2_grid); cfg.blockDim = dim3(kC2Threads); cfg.dynamicSmemBytes = h->c2_smem_bytes; cfg.stream = h->stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;      // (the cluster shape is the kernel's own attribute)
        cfg.attrs = at; cfg.numAttrs = 1;
        const int scheme = ea.world > 1 ? (x.rsag ? 2 : 1) : 0;
        const bool waves = n_pairs > h->c2_grid / 2;
        if (waves && h->env_identity && h->series_len * (int64_t)sizeof(float) > 0) {
            QP_TRY(c2_window_prepare(h));
            x.win_tma = 1;
        }
        auto* kern = x.split ? (scheme == 2 ? episode_c2_kernel<2, true, false> : scheme == 1 ? episode_c2_kernel<1, true, false> : episode_c2_kernel<0, true, false>)
                     : waves ? (scheme == 2 ? episode_c2_kernel<2, false, true> : scheme == 1 ? episode_c2_kernel<1, false, true> : episode_c2_kernel<0, false, true>)
                             : (scheme == 2 ? episode_c2_kernel<2, false, false> : scheme == 1 ? episode_c2_kernel<1, false, false> : episode_c2_kernel<0, false, false>);
        QP_CUDA(cudaLaunchKernelEx(&cfg, kern, h->tm_c2_hi, h->tm_c2_lo, h->tm_c2_part, x.win_tma ? h->tm_c2_win : h->tm_c2_part, x));
        h->launches++;
        return QP_OK;
    }
    if (ep_tc_usable(h) && apply_in_kernel) {
        EpTcArgs x{};
        x.ea = ea;
        x.w1T_hi = h->et_w_hi.p; x.w1T_lo = h->et_w_lo.p;
        x.Kp = h->et_Kp; x.nkb = h->et_nkb; x.Np = h->et_Np; x.halves = h->et_halves;
        if (h->want_cta_ns) {        // the stamps share the dbg buffer: [n_steps][grid][4] instead of [n_steps][kDbgSlots]
            QP_TRY(h->dbg.ensure((size_t)n_steps * h->num_sms * 4));
            x.ea.dbg = nullptr;
            x.cta_ns = h->dbg.p;
        }
        QP_TRY(h->et_published.ensure(2));
        x.published = h->et_published.p;
        x.p_pad = (int64_t)p2p_pad(h);
        for (int r = 0; r < ea.world && ea.world > 1; ++r) {
            x.pdata[r] = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(h->p2p_peer_block[r]) + p2p_tc_offset(h));
        }
        if (h->tm_et_part_ptr != h->part1.p || h->tm_et_part_tiles != n_tiles) {
            QP_TRY(make_tmap3_store(&h->tm_et_part, h->part1.p, (uint64_t)h->hp.H, (uint64_t)h->hp.wn, (uint64_t)n_tiles,
                                    (uint64_t)h->hp.H * sizeof(float), (uint64_t)h->P1 * sizeof(float)));
            h->tm_et_part_ptr = h->part1.p; h->tm_et_part_tiles = n_tiles;
        }
        void* targs[] = {&h->tm_et_hi, &h->tm_et_lo, &h->tm_et_part, &x};
        QP_CUDA(cudaLaunchCooperativeKernel((const void*)episode_tc_kernel, dim3(h->num_sms), dim3(kEtThreads),
                                            targs, h->et_smem, h->stream));
        h->launches++;
        return QP_OK;
    }
    void* args[] = {&ea};
    QP_CUDA(cudaLaunchCooperativeKernel((const void*)episode_kernel, dim3(h->num_sms), dim3(h->geo.ep_threads),
                                        args, h->geo.smem_ep, h->stream));
    h->launches++;
    return QP_OK;
}

// Several GPUs, episodes fit one tile per SM: each env-step is the persistent kernel's phase A as ONE
// launch (forward, select, book-keeping and the dW1 contraction fused), then sum of partials,
// all-reduce of the gradient, AdaGrad.
bool use_fused_step(const qp_handle* h) {
    return !h->wide && h->allow_fused && h->coop_launch && h->world > 1 && !use_fused(h) && h->geo.ep_ok &&
           h->n_envs <= (int64_t)kTileEnvs * h->num_sms;
}

int enqueue_env_step_fused_dist(qp_handle* h, const TraceDev* trace, int32_t trace_step0, int32_t step_offset) {
    QP_TRY(enqueue_env_steps_fused(h, 1, trace, trace_step0, false, step_offset));
    const int n_tiles = (int)(round_up(h->n_envs, kTileEnvs) / kTileEnvs);
    QP_TRY(launch_apply(h, n_tiles, n_tiles, 1, 0, 0));
    QP_NCCL(g_nccl.AllReduce(h->grad.p, h->grad.p, (size_t)h->P, kNcclFloat32, kNcclSum, h->comm, h->stream));
    QP_TRY(launch_apply(h, 0, 0, 0, 1, 1));
    return QP_OK;
}

int check_env_steps(qp_handle* h, int32_t n_steps) {
    if (!h) return fail(QP_EINVAL, "null handle");
    if (!h->has_env) return fail(QP_ESTATE, "no episodes: call qp_env_create first");
    if (n_steps < 0) return fail(QP_EINVAL, "n_steps < 0");
    const int64_t total = h->series_len - h->hp.wn;
    if (h->steps_done + n_steps > total)
        return fail(QP_EINVAL, "episode has " + std::to_string(total) + " steps, " +
                                   std::to_string(h->steps_done) + " done, asked for " + std::to_string(n_steps));
    return QP_OK;
}

}  // namespace

extern "C" {

const char* qp_last_error(void) { return g_err.c_str(); }
const char* qp_version(void) { return "qpolicy-b200 0.1 (sm_100a)"; }

int qp_config_default(qp_config* cfg) {
    if (!cfg) return fail(QP_EINVAL, "null cfg");
    cfg->window = 201; cfg->hidden = 200;
    cfg->epsilon = 0.9f; cfg->gamma = 0.001f; cfg->lr = 0.01f; cfg->bias = 0.1f;
    cfg->step_div = 1000.0f; cfg->adagrad_init_acc = 1e-8f;
    cfg->bookkeeping = QP_BOOK_REFERENCE; cfg->device = 0; cfg->seed = 42;
    return QP_OK;
}

int qp_create(const qp_config* cfg, qp_handle** out) {
    if (!cfg || !out) return fail(QP_EINVAL, "null argument");
    *out = nullptr;
    if (cfg->window < 1 || cfg->hidden < 4) return fail(QP_EINVAL, "window >= 1 and hidden >= 4 required");
    if (cfg->hidden % 4) return fail(QP_EINVAL, "hidden must be a multiple of 4");
    if (cfg->bookkeeping != QP_BOOK_REFERENCE && cfg->bookkeeping != QP_BOOK_RLPY)
        return fail(QP_EINVAL, "bookkeeping must be QP_BOOK_REFERENCE or QP_BOOK_RLPY");
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0)
        return fail(QP_ECUDA, std::string("no CUDA device: ") + cudaGetErrorString(ce) +
                                  " (this library has no CPU path)");
    if (cfg->device < 0 || cfg->device >= ndev) return fail(QP_EINVAL, "device ordinal out of range");

    qp_handle* h = new qp_handle();
    h->cfg = *cfg;
    h->device = cfg->device;
    Hyper& hp = h->hp;
    hp.wn = cfg->window; hp.D = cfg->window + 2; hp.H = cfg->hidden;
    hp.Hs = (int)round_up(cfg->hidden, kColsPerWarp);
    hp.epsilon = cfg->epsilon; hp.gamma = cfg->gamma; hp.lr = cfg->lr; hp.bias = cfg->bias;
    hp.step_div = cfg->step_div; hp.bookkeeping = cfg->bookkeeping; hp.seed = cfg->seed;
    h->P1 = (int64_t)hp.D * hp.H;
    h->P = h->P1 + 3LL * hp.H;

    auto bail = [&](int code) { qp_destroy(h); return code; };
    if (cudaSetDevice(h->device) != cudaSuccess) return bail(fail(QP_ECUDA, "cudaSetDevice failed"));
    cudaDeviceProp prop{};
    if (cudaGetDeviceProperties(&prop, h->device) != cudaSuccess) return bail(fail(QP_ECUDA, "cudaGetDeviceProperties failed"));
    h->num_sms = prop.multiProcessorCount;
    h->max_smem = (int)prop.sharedMemPerBlockOptin;

    // launch geometry and shape limits of the register-tiled fp32 path
    h->tile_ok = make_geometry(hp, h->max_smem, &h->geo) == QP_OK;
    if (!h->tile_ok) {
        // W1 does not fit in shared memory: organise the step around GEMMs (qp_wide.cuh)
        if (hp.H % 128 != 0 || hp.H / 32 > kWideMaxColsPerLane)
            return bail(fail(QP_EUNSUPPORTED, "shape outside both paths: the per-tile path needs W1 resident in shared "
                                              "memory (hidden <= 240, (window+2)*hidden*4 + staging <= " +
                                                  std::to_string(h->max_smem) + " bytes); the GEMM path needs hidden to be a "
                                              "multiple of 128 and <= 2048"));
        h->wide = true;
        h->geo.ep_ok = false;
    }
#define QP_BAIL_CUDA(expr)                                                                        \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess) return bail(fail(QP_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(_e))); \
    } while (0)
    // The attribute is per function and per device, not per handle: always raise it to the device
    // maximum so that handles of different shapes can coexist.
    QP_BAIL_CUDA(cudaFuncSetAttribute(step_kernel<kEnvStep>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem));
    QP_BAIL_CUDA(cudaFuncSetAttribute(step_kernel<kMsgSelect>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem));
    QP_BAIL_CUDA(cudaFuncSetAttribute(step_kernel<kMsgUpdate>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem));
    QP_BAIL_CUDA(cudaFuncSetAttribute(dw1_kernel<kEnvStep>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem));
    QP_BAIL_CUDA(cudaFuncSetAttribute(dw1_kernel<kMsgUpdate>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem));
    QP_BAIL_CUDA(cudaFuncSetAttribute(dw1_big_kernel<kEnvStep>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem));
    QP_BAIL_CUDA(cudaFuncSetAttribute(dw1_big_kernel<kMsgUpdate>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem));
    QP_BAIL_CUDA(cudaFuncSetAttribute(episode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem));
    QP_BAIL_CUDA(cudaFuncSetAttribute(episode_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem));
    QP_BAIL_CUDA(cudaFuncSetAttribute(wide_row_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem));
    QP_BAIL_CUDA(cudaFuncSetAttribute(wide_row_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem));
    QP_BAIL_CUDA(cudaFuncSetAttribute(wide_row_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem));
    {
        int coop = 0;
        QP_BAIL_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->device));
        h->coop_launch = coop != 0;
    }
    QP_BAIL_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    for (auto& e : h->ev) QP_BAIL_CUDA(cudaEventCreate(&e));
#undef QP_BAIL_CUDA
    if (h->params.ensure(h->P) || h->acc.ensure(h->P) || h->grad.ensure(h->P) || h->d_step.ensure(1) || h->flags4.ensure(8))
        return bail(QP_ECUDA);
    if (cudaHostAlloc(reinterpret_cast<void**>(&h->host_flag), sizeof(int32_t), cudaHostAllocDefault) != cudaSuccess)
        return bail(fail(QP_ECUDA, "cudaHostAlloc failed"));

    std::vector<float> w(h->P);
    gen_normal(cfg->seed, h->P, w.data());
    int st = qp_set_weights(h, w.data(), nullptr);
    if (st != QP_OK) return bail(st);
    std::vector<float> a(h->P, cfg->adagrad_init_acc);
    st = qp_set_weights(h, nullptr, a.data());
    if (st != QP_OK) return bail(st);
    if (cudaMemsetAsync(h->grad.p, 0, h->P * sizeof(float), h->stream) != cudaSuccess ||
        cudaMemsetAsync(h->d_step.p, 0, sizeof(int32_t), h->stream) != cudaSuccess ||
        cudaStreamSynchronize(h->stream) != cudaSuccess)
        return bail(fail(QP_ECUDA, "initial memset failed"));
    *out = h;
    return QP_OK;
}

int qp_destroy(qp_handle* h) {
    if (!h) return QP_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(h->comm);
    if (!h->p2p_external) {
        for (int r = 0; r < 8; ++r)
            if (h->p2p_peer_block[r] && r != h->rank) cudaIpcCloseMemHandle(h->p2p_peer_block[r]);
        if (h->p2p_block) cudaFree(h->p2p_block);
    }
    h->params.release(); h->acc.release(); h->grad.release();
    h->dz1.release(); h->xextra.release(); h->part1.release(); h->part2.release();
    h->m_states.release(); h->m_next.release(); h->m_rewards.release(); h->m_steps.release();
    h->m_q.release(); h->m_ids.release(); h->m_actions.release();
    h->wZ1.release(); h->wdZ1.release(); h->wdz2.release(); h->wkn.release();
    h->et_w_hi.release(); h->et_w_lo.release(); h->et_published.release();
    for (int g = 0; g < 2; ++g)
        for (qp_handle::FrontQueue* fq : {&h->q_sel[g], &h->q_upd[g]}) {
            if (fq->states) cudaFreeHost(fq->states);
            if (fq->scalars) cudaFreeHost(fq->scalars);
            if (fq->next) cudaFreeHost(fq->next);
            if (fq->ids) cudaFreeHost(fq->ids);
        }
    for (int j = 0; j < 4; ++j) { h->p_hi[j].release(); h->p_lo[j].release(); }
    h->pT_hi.release(); h->pT_lo.release(); h->xe_hi.release(); h->xe_lo.release();
    h->w1T_hi.release(); h->w1T_lo.release(); h->dzT_hi.release(); h->dzT_lo.release();
    h->prices.release(); h->series_of_env.release(); h->shares.release();
    h->budget.release(); h->share_value.release(); h->portfolio.release(); h->blocksum.release();
    h->d_step.release(); h->dbg.release(); h->flags4.release();
    h->fan_chunks.release(); h->fan_zw.release(); h->fan_part.release(); h->c2_pT.release();
    if (h->host_flag) cudaFreeHost(h->host_flag);
    for (auto& e : h->ev) if (e) cudaEventDestroy(e);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return QP_OK;
}

int64_t qp_input_dim(const qp_handle* h) { return h ? h->hp.D : 0; }
int64_t qp_param_count(const qp_handle* h) { return h ? h->P : 0; }
void* qp_stream(qp_handle* h) { return h ? (void*)h->stream : nullptr; }
int64_t qp_launch_count(const qp_handle* h) { return h ? h->launches : 0; }

int qp_set_weights(qp_handle* h, const float* params, const float* acc) {
    if (!h) return fail(QP_EINVAL, "null handle");
    QP_TRY(use_device(h));
    if (params) QP_CUDA(cudaMemcpyAsync(h->params.p, params, h->P * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    if (acc) QP_CUDA(cudaMemcpyAsync(h->acc.p, acc, h->P * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    QP_CUDA(cudaStreamSynchronize(h->stream));
    return QP_OK;
}

int qp_get_weights(qp_handle* h, float* params, float* acc) {
    if (!h) return fail(QP_EINVAL, "null handle");
    QP_TRY(use_device(h));
    if (params) QP_CUDA(cudaMemcpyAsync(params, h->params.p, h->P * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    if (acc) QP_CUDA(cudaMemcpyAsync(acc, h->acc.p, h->P * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    QP_CUDA(cudaStreamSynchronize(h->stream));
    return QP_OK;
}

int qp_get_last_grad(qp_handle* h, float* grad) {
    if (!h || !grad) return fail(QP_EINVAL, "null argument");
    QP_TRY(use_device(h));
    QP_CUDA(cudaMemcpyAsync(grad, h->grad.p, h->P * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    QP_CUDA(cudaStreamSynchronize(h->stream));
    return QP_OK;
}

// SelectionAction / UpdateQ on the GEMM-organised path (policies whose W1 does not fit in shared memory): the rows the
// caller hands over take the place of the price windows.  G1 (fp32 FFMA GEMM over the window part of the rows), the row
// kernel in its message form, and for UpdateQ E2, G2 over all D inputs and the AdaGrad step.  msg = 1 select, 2 update.
int wide_message(qp_handle* h, int msg, int64_t B, bool have_ids) {
    const Hyper& hp = h->hp;
    const int64_t Bpad = round_up(B, kTileEnvs);
    const int tiles_m = (hp.D + kWgBM - 1) / kWgBM, tiles_n = (hp.H + kWgBN - 1) / kWgBN;
    const int S = (int)std::max<int64_t>(1, std::min<int64_t>({(int64_t)32, B / 256 > 0 ? B / 256 : 1,
                                                               (h->num_sms * 6 + tiles_m * tiles_n - 1) / (tiles_m * tiles_n)}));
    const int colblocks = (hp.H + 255) / 256;
    const int S2 = (int)std::max<int64_t>(1, std::min<int64_t>(B / 64 > 0 ? B / 64 : 1, (h->num_sms * 2) / colblocks));
    const int64_t rows = msg == 1 ? B : 2 * B;
    QP_TRY(h->wZ1.ensure((size_t)rows * hp.H));
    WideArgs a{};
    a.hp = hp; a.params = h->params.p; a.d_step = h->d_step.p;
    a.B = B; a.Bpad = Bpad; a.Z1 = h->wZ1.p; a.msg = msg;
    a.m_states = h->m_states.p; a.m_steps = h->m_steps.p; a.m_ids = have_ids ? h->m_ids.p : nullptr;
    a.m_actions = h->m_actions.p; a.m_q = msg == 1 ? h->m_q.p : nullptr;
    if (msg == 2) {
        QP_TRY(h->wdZ1.ensure((size_t)B * hp.H));
        QP_TRY(h->wdz2.ensure(B));
        QP_TRY(h->wkn.ensure(B));
        QP_TRY(h->xextra.ensure((size_t)Bpad * 2));
        QP_TRY(h->part1.ensure((size_t)S * h->P1));
        QP_TRY(h->part2.ensure((size_t)S2 * 3 * hp.H));
        a.m_next = h->m_next.p; a.m_rewards = h->m_rewards.p;
        a.dZ1 = h->wdZ1.p; a.xextra = h->xextra.p; a.kn = h->wkn.p; a.dz2 = h->wdz2.p;
        a.part1 = h->part1.p; a.part2 = h->part2.p; a.n_split = S; a.n_split2 = S2;
    }
    wide_gemm_kernel<0><<<dim3((unsigned)((rows + kWgBM - 1) / kWgBM), tiles_n), kWgThreads, 0, h->stream>>>(a);
    QP_CUDA(cudaGetLastError());
    const int row_blocks = (int)std::min<int64_t>((B + 7) / 8, (int64_t)h->num_sms * 4);
    if (msg == 1) {
        wide_row_kernel<1><<<row_blocks, 256, (size_t)hp.H * 6 * sizeof(float), h->stream>>>(a);
        QP_CUDA(cudaGetLastError());
        h->launches += 2;
        return QP_OK;
    }
    wide_row_kernel<2><<<row_blocks, 256, (size_t)hp.H * 6 * sizeof(float), h->stream>>>(a);
    QP_CUDA(cudaGetLastError());
    wide_dw2_kernel<<<dim3(colblocks, S2), 256, 0, h->stream>>>(a);
    QP_CUDA(cudaGetLastError());
    wide_gemm_kernel<1><<<dim3(tiles_m, tiles_n, S), kWgThreads, 0, h->stream>>>(a);
    QP_CUDA(cudaGetLastError());
    h->launches += 4;
    QP_TRY(launch_apply(h, S, S2, 1, 1, 0));
    return QP_OK;
}

int qp_select_action(qp_handle* h, const float* states, int64_t state_len, const float* steps,
                     const int64_t* env_ids, int32_t* actions, float* q_out, int64_t B) {
    if (!h || !states || !steps || !actions) return fail(QP_EINVAL, "null argument");
    if (state_len != h->hp.D)  // QDecisionPolicyActor.scala:56-57
        return fail(QP_EINVAL, "SelectionAction: tensorflow input size(" + std::to_string(h->hp.D) +
                                   ") and state(" + std::to_string(state_len) + ") size do not match");
    if (B < 0) return fail(QP_EINVAL, "B < 0");
    if (B == 0) return QP_OK;
    QP_TRY(use_device(h));
    const int D = h->hp.D;
    QP_TRY(h->m_states.ensure((size_t)B * D));
    QP_TRY(h->m_steps.ensure(B));
    QP_TRY(h->m_actions.ensure(B));
    QP_TRY(h->m_q.ensure((size_t)B * 3));
    if (env_ids) QP_TRY(h->m_ids.ensure(B));
    QP_CUDA(cudaMemcpyAsync(h->m_states.p, states, (size_t)B * D * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    QP_CUDA(cudaMemcpyAsync(h->m_steps.p, steps, B * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    if (env_ids) QP_CUDA(cudaMemcpyAsync(h->m_ids.p, env_ids, B * sizeof(int64_t), cudaMemcpyHostToDevice, h->stream));
    if (h->wide) {
        QP_TRY(wide_message(h, 1, B, env_ids != nullptr));
    } else {
        StepArgs s{};
        s.hp = h->hp; s.params = h->params.p;
        s.states = h->m_states.p; s.steps = h->m_steps.p; s.env_ids = env_ids ? h->m_ids.p : nullptr;
        s.actions_out = h->m_actions.p; s.q_out = h->m_q.p; s.B = B; s.Bpad = round_up(B, kTileEnvs);
        const int grid = (int)std::min<int64_t>(s.Bpad / kTileEnvs, h->num_sms);
        QP_TRY(launch_step<kMsgSelect>(h, s, grid));
    }
    QP_CUDA(cudaMemcpyAsync(actions, h->m_actions.p, B * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    if (q_out) QP_CUDA(cudaMemcpyAsync(q_out, h->m_q.p, (size_t)B * 3 * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    QP_CUDA(cudaStreamSynchronize(h->stream));
    return QP_OK;
}

int qp_update_q(qp_handle* h, const float* states, int64_t state_len, const float* rewards,
                const float* next_states, int64_t next_state_len, int64_t B) {
    if (!h || !states || !rewards || !next_states) return fail(QP_EINVAL, "null argument");
    if (state_len != h->hp.D || next_state_len != h->hp.D)  // QDecisionPolicyActor.scala:64-65
        return fail(QP_EINVAL, "update q: tensorflow input size(" + std::to_string(h->hp.D) +
                                   ") and state(or nextState) size(" + std::to_string(state_len) + ", " +
                                   std::to_string(next_state_len) + ") do not match");
    if (B < 0) return fail(QP_EINVAL, "B < 0");
    if (B == 0) return QP_OK;
    QP_TRY(use_device(h));
    const int D = h->hp.D;
    int g1 = 0, g2 = 0;
    if (!h->wide) QP_TRY(ensure_scratch(h, B, &g1, &g2));
    QP_TRY(h->m_states.ensure((size_t)B * D));
    QP_TRY(h->m_next.ensure((size_t)B * D));
    QP_TRY(h->m_rewards.ensure(B));
    QP_CUDA(cudaMemcpyAsync(h->m_states.p, states, (size_t)B * D * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    QP_CUDA(cudaMemcpyAsync(h->m_next.p, next_states, (size_t)B * D * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    QP_CUDA(cudaMemcpyAsync(h->m_rewards.p, rewards, B * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    if (h->wide) {
        QP_TRY(wide_message(h, 2, B, false));
    } else {
        StepArgs s{};
        s.hp = h->hp; s.params = h->params.p;
        s.states = h->m_states.p; s.next_states = h->m_next.p; s.rewards = h->m_rewards.p;
        s.B = B; s.Bpad = h->Bpad; s.dz1 = h->dz1.p; s.xextra = h->xextra.p; s.dw2_part = h->part2.p;
        QP_TRY(launch_step<kMsgUpdate>(h, s, g1));
        Dw1Args d{};
        d.hp = h->hp; d.states = h->m_states.p; d.Bpad = h->Bpad; d.dz1 = h->dz1.p; d.B = B; d.part = h->part1.p;
        d.n_rt = h->geo.n_rt; d.cg_max = h->geo.cg_max;
        QP_TRY(launch_dw1<kMsgUpdate>(h, d, g2));
        QP_TRY(launch_apply(h, g2, g1, 1, 1, 0));
    }
    QP_CUDA(cudaStreamSynchronize(h->stream));
    // if (iteration % 500 == 0 && iteration != 0) saveSnapshot(...)  — QDecisionPolicyActor.scala:74;
    // the reference's saveSnapshot is an empty stub (:91-93), here it writes a checkpoint if a path is set
    // The trigger is the reference's (the iteration number this message was handled with); the checkpoint header
    // stores the number of updates the saved weights CONTAIN, so that a resumed policy neither repeats the
    // snapshot nor lags by one.
    const int64_t handled_with = h->iteration;
    h->iteration = handled_with + 1;     // context.become(iterateUpdate(session, iteration + 1)) (:76)
    if (!h->snapshot_path.empty() && h->snapshot_every > 0 && handled_with != 0 &&
        handled_with % h->snapshot_every == 0)
        QP_TRY(qp_save_checkpoint(h, h->snapshot_path.c_str()));
    return QP_OK;
}

int qp_env_create(qp_handle* h, const float* prices, int64_t series_len, int64_t n_series,
                  const int32_t* series_of_env, int64_t n_envs, int64_t env_id_base,
                  double budget0, int32_t shares0) {
    if (!h || !prices || !series_of_env) return fail(QP_EINVAL, "null argument");
    if (series_len <= h->hp.wn)  // TrainerChildActor.scala:65, :70
        return fail(QP_EINVAL, "Stock price count should be more than Tensorflow input nodes");
    if (n_series < 1 || n_envs < 1) return fail(QP_EINVAL, "n_series and n_envs must be >= 1");
    if (series_len > INT32_MAX) return fail(QP_EINVAL, "series too long");
    for (int64_t e = 0; e < n_envs; ++e)
        if (series_of_env[e] < 0 || series_of_env[e] >= n_series)
            return fail(QP_EINVAL, "series_of_env[" + std::to_string(e) + "] out of range");
    QP_TRY(use_device(h));
    QP_TRY(h->prices.ensure((size_t)n_series * series_len));
    QP_TRY(h->series_of_env.ensure(n_envs));
    QP_TRY(h->budget.ensure(n_envs));
    QP_TRY(h->shares.ensure(n_envs));
    QP_TRY(h->share_value.ensure(n_envs));
    QP_TRY(h->portfolio.ensure(n_envs));
    QP_TRY(h->blocksum.ensure(1024));
    QP_CUDA(cudaMemcpyAsync(h->prices.p, prices, (size_t)n_series * series_len * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    QP_CUDA(cudaMemcpyAsync(h->series_of_env.p, series_of_env, n_envs * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    h->series_len = series_len; h->n_series = n_series; h->n_envs = n_envs; h->env_id_base = env_id_base;
    h->B0 = budget0; h->N0 = shares0;
    h->has_env = true;
    h->tc_ready = false;
    h->c2_win_ready = false;
    h->env_identity = n_series == n_envs;
    for (int64_t e = 0; h->env_identity && e < n_envs; ++e) h->env_identity = series_of_env[e] == e;
    {   // fan-out path: are the episodes grouped by series (series_of_env non-decreasing)?  Then cut every series'
        // run of episodes into chunks of at most kFanChunk.
        h->fan_ok = n_envs < ((int64_t)1 << 31);
        for (int64_t e = 1; h->fan_ok && e < n_envs; ++e) h->fan_ok = series_of_env[e] >= series_of_env[e - 1];
        h->fan_chunks_host.clear();
        h->fan_n_chunks = 0;
        if (h->fan_ok) {
            for (int64_t e = 0; e < n_envs;) {
                int64_t end = e + 1;
                while (end < n_envs && end - e < kFanChunk && series_of_env[end] == series_of_env[e]) ++end;
                h->fan_chunks_host.insert(h->fan_chunks_host.end(), {series_of_env[e], (int32_t)e, (int32_t)(end - e), 0});
                e = end;
            }
            h->fan_n_chunks = (int)(h->fan_chunks_host.size() / 4);
            QP_TRY(h->fan_chunks.ensure(h->fan_chunks_host.size()));
            QP_CUDA(cudaMemcpyAsync(h->fan_chunks.p, h->fan_chunks_host.data(), h->fan_chunks_host.size() * sizeof(int32_t),
                                    cudaMemcpyHostToDevice, h->stream));      // (qp_env_reset below synchronises)
        }
    }
    bool flag_pending = false;
    QP_TRY(ep_tc_prepare(h, &flag_pending));
    if (h->wide || (h->hp.H % 128 == 0 && h->hp.H / 32 <= kWideMaxColsPerLane)) {
        bool identity = n_series == n_envs;
        for (int64_t e = 0; identity && e < n_envs; ++e) identity = series_of_env[e] == e;
        // the tensor-core GEMMs read the price windows straight from the price matrix with TMA, which
        // needs episode e <-> series e and 16-byte row pitches
        if (identity && series_len % 4 == 0) QP_TRY(wide_prepare_tc(h));
    }
    QP_TRY(qp_env_reset(h));                       // (synchronises the stream)
    return ep_tc_finish(h, flag_pending);
}

int qp_env_reset(qp_handle* h) {
    if (!h) return fail(QP_EINVAL, "null handle");
    if (!h->has_env) return fail(QP_ESTATE, "no episodes: call qp_env_create first");
    QP_TRY(use_device(h));
    const int blocks = (int)((h->n_envs + 255) / 256);
    env_init_kernel<<<blocks, 256, 0, h->stream>>>(h->budget.p, h->shares.p, h->share_value.p, h->n_envs, h->B0, h->N0);
    h->launches++;
    QP_CUDA(cudaGetLastError());
    QP_CUDA(cudaMemsetAsync(h->d_step.p, 0, sizeof(int32_t), h->stream));
    QP_CUDA(cudaStreamSynchronize(h->stream));
    h->steps_done = 0;
    return QP_OK;
}

int64_t qp_env_steps_total(const qp_handle* h) { return (h && h->has_env) ? h->series_len - h->hp.wn : 0; }
int64_t qp_env_steps_done(const qp_handle* h) { return h ? h->steps_done : 0; }

int qp_env_step_async(qp_handle* h, int32_t n_steps) {
    QP_TRY(check_env_steps(h, n_steps));
    QP_TRY(use_device(h));
    if (h->wide) {
        for (int32_t i = 0; i < n_steps; ++i) QP_TRY(enqueue_env_step_wide(h, nullptr, 0, h->steps_done + i));
    } else if (use_fanout(h)) {
        for (int32_t i = 0; i < n_steps; ++i) QP_TRY(enqueue_env_step_fanout(h, nullptr, 0));
    } else if (use_fused(h)) {
        QP_TRY(enqueue_env_steps_fused(h, n_steps, nullptr, 0));
    } else if (use_fused_step(h)) {
        for (int32_t i = 0; i < n_steps; ++i) QP_TRY(enqueue_env_step_fused_dist(h, nullptr, 0, i));
    } else {
        for (int32_t i = 0; i < n_steps; ++i) QP_TRY(enqueue_env_step(h, nullptr, 0, false));
    }
    h->steps_done += n_steps;
    return QP_OK;
}

int qp_sync(qp_handle* h) {
    if (!h) return fail(QP_EINVAL, "null handle");
    QP_TRY(use_device(h));
    QP_CUDA(cudaStreamSynchronize(h->stream));
    return QP_OK;
}

int qp_env_step(qp_handle* h, int32_t n_steps) {
    QP_TRY(qp_env_step_async(h, n_steps));
    return qp_sync(h);
}

int qp_env_step_timed(qp_handle* h, int32_t n_steps, float* ms_sum) {
    if (!ms_sum) return fail(QP_EINVAL, "null ms_sum");
    QP_TRY(check_env_steps(h, n_steps));
    QP_TRY(use_device(h));
    for (int k = 0; k < QP_N_KERNELS; ++k) ms_sum[k] = 0.0f;
    if (h->wide) {        // five kernels per step; the whole batch of steps is timed into slot QP_KERNEL_STEP
        QP_CUDA(cudaEventRecord(h->ev[0], h->stream));
        for (int32_t i = 0; i < n_steps; ++i) QP_TRY(enqueue_env_step_wide(h, nullptr, 0, h->steps_done + i));
        QP_CUDA(cudaEventRecord(h->ev[1], h->stream));
        QP_CUDA(cudaEventSynchronize(h->ev[1]));
        QP_CUDA(cudaEventElapsedTime(&ms_sum[0], h->ev[0], h->ev[1]));
        h->steps_done += n_steps;
        return QP_OK;
    }
    const bool fan = use_fanout(h);
    if (!fan && use_fused_step(h)) {   // fused step kernel + sum + all-reduce + AdaGrad per step, timed as a whole
        QP_CUDA(cudaEventRecord(h->ev[0], h->stream));
        for (int32_t i = 0; i < n_steps; ++i) QP_TRY(enqueue_env_step_fused_dist(h, nullptr, 0, i));
        QP_CUDA(cudaEventRecord(h->ev[1], h->stream));
        QP_CUDA(cudaEventSynchronize(h->ev[1]));
        QP_CUDA(cudaEventElapsedTime(&ms_sum[0], h->ev[0], h->ev[1]));
        h->steps_done += n_steps;
        return QP_OK;
    }
    if (!fan && use_fused(h)) {   // one persistent kernel: its whole duration goes to slot QP_KERNEL_STEP
        QP_CUDA(cudaEventRecord(h->ev[0], h->stream));
        QP_TRY(enqueue_env_steps_fused(h, n_steps, nullptr, 0));
        QP_CUDA(cudaEventRecord(h->ev[1], h->stream));
        QP_CUDA(cudaEventSynchronize(h->ev[1]));
        QP_CUDA(cudaEventElapsedTime(&ms_sum[0], h->ev[0], h->ev[1]));
        h->steps_done += n_steps;
        return QP_OK;
    }
    // Events are recorded for a batch of steps without any host synchronisation in between, so the
    // kernels run back to back as they do in qp_env_step; each interval is one kernel plus the
    // launch gap that follows it.
    constexpr int kBatch = 32;
    std::vector<cudaEvent_t> evs((size_t)kBatch * 4, nullptr);
    for (auto& e : evs) QP_CUDA(cudaEventCreate(&e));
    int st = QP_OK;
    for (int32_t done = 0; done < n_steps && st == QP_OK; done += kBatch) {
        const int nb = std::min<int32_t>(kBatch, n_steps - done);
        for (int i = 0; i < nb && st == QP_OK; ++i) {
            for (int k = 0; k < 4; ++k) h->ev[k] = evs[(size_t)i * 4 + k];
            // (fan-out path: slots = F1 | F2 | F3 + AdaGrad)
            st = use_fanout(h) ? enqueue_env_step_fanout(h, nullptr, 0, true) : enqueue_env_step(h, nullptr, 0, true);
        }
        if (st == QP_OK && cudaStreamSynchronize(h->stream) != cudaSuccess) st = fail(QP_ECUDA, "sync failed");
        for (int i = 0; i < nb && st == QP_OK; ++i)
            for (int k = 0; k < QP_N_KERNELS; ++k) {
                float ms = 0.0f;
                if (cudaEventElapsedTime(&ms, evs[(size_t)i * 4 + k], evs[(size_t)i * 4 + k + 1]) != cudaSuccess)
                    st = fail(QP_ECUDA, "cudaEventElapsedTime failed");
                ms_sum[k] += ms;
            }
        if (st == QP_OK) h->steps_done += nb;
    }
    for (int k = 0; k < 4; ++k) h->ev[k] = nullptr;
    for (auto& e : evs) cudaEventDestroy(e);
    for (auto& e : h->ev) if (cudaEventCreate(&e) != cudaSuccess && st == QP_OK) st = fail(QP_ECUDA, "cudaEventCreate failed");
    return st;
}

int qp_env_step_trace(qp_handle* h, int32_t n_steps, const qp_trace* out) {
    if (!out) return fail(QP_EINVAL, "null trace");
    QP_TRY(check_env_steps(h, n_steps));
    if (n_steps == 0) return QP_OK;
    QP_TRY(use_device(h));
    const size_t n = (size_t)n_steps * h->n_envs;
    DevBuf<int32_t> asel, aeff, expl, kn, shr;
    DevBuf<float> rew, qs, qn, tgt;
    DevBuf<double> bud;
    TraceDev t{};
    int st = QP_OK;
    auto rel = [&]() { asel.release(); aeff.release(); expl.release(); kn.release(); shr.release();
                       rew.release(); qs.release(); qn.release(); tgt.release(); bud.release(); };
#define QP_TR(buf, field, host, mult)                                  \
    if (st == QP_OK && out->host) { st = buf.ensure(n * mult); t.field = buf.p; }
    QP_TR(asel, action_sel, action_sel, 1) QP_TR(aeff, action_eff, action_eff, 1)
    QP_TR(expl, explored, explored, 1) QP_TR(kn, k_next, k_next, 1) QP_TR(shr, shares, shares, 1)
    QP_TR(rew, reward, reward, 1) QP_TR(qs, q_s, q_s, 3) QP_TR(qn, q_n, q_n, 3)
    QP_TR(tgt, target, target, 1) QP_TR(bud, budget, budget, 1)
#undef QP_TR
    const int32_t step0 = (int32_t)h->steps_done;
    if (h->wide) {
        for (int32_t i = 0; i < n_steps && st == QP_OK; ++i) st = enqueue_env_step_wide(h, &t, step0, h->steps_done + i);
    } else if (use_fanout(h)) {
        for (int32_t i = 0; i < n_steps && st == QP_OK; ++i) st = enqueue_env_step_fanout(h, &t, step0);
    } else if (use_fused(h)) {
        st = enqueue_env_steps_fused(h, n_steps, &t, step0);
    } else if (use_fused_step(h)) {
        for (int32_t i = 0; i < n_steps && st == QP_OK; ++i) st = enqueue_env_step_fused_dist(h, &t, step0, i);
    } else {
        for (int32_t i = 0; i < n_steps && st == QP_OK; ++i) st = enqueue_env_step(h, &t, step0, false);
    }
    if (st == QP_OK) h->steps_done += n_steps;
    auto d2h = [&](void* dst, const void* src, size_t bytes) {
        if (st == QP_OK && dst && cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, h->stream) != cudaSuccess)
            st = fail(QP_ECUDA, std::string("trace copy failed: ") + cudaGetErrorString(cudaGetLastError()));
    };
    d2h(out->action_sel, asel.p, n * 4); d2h(out->action_eff, aeff.p, n * 4); d2h(out->explored, expl.p, n * 4);
    d2h(out->k_next, kn.p, n * 4); d2h(out->shares, shr.p, n * 4); d2h(out->reward, rew.p, n * 4);
    d2h(out->q_s, qs.p, n * 12); d2h(out->q_n, qn.p, n * 12); d2h(out->target, tgt.p, n * 4);
    d2h(out->budget, bud.p, n * 8);
    if (cudaStreamSynchronize(h->stream) != cudaSuccess && st == QP_OK) st = fail(QP_ECUDA, "sync failed after trace");
    rel();
    return st;
}

static int compute_portfolio(qp_handle* h) {
    const int blocks = (int)((h->n_envs + 255) / 256);
    portfolio_kernel<<<blocks, 256, 0, h->stream>>>(h->budget.p, h->shares.p, h->share_value.p, h->n_envs, h->portfolio.p);
    h->launches++;
    QP_CUDA(cudaGetLastError());
    return QP_OK;
}

int qp_env_get_portfolio(qp_handle* h, double* out) {
    if (!h || !out) return fail(QP_EINVAL, "null argument");
    if (!h->has_env) return fail(QP_ESTATE, "no episodes: call qp_env_create first");
    QP_TRY(use_device(h));
    QP_TRY(compute_portfolio(h));
    QP_CUDA(cudaMemcpyAsync(out, h->portfolio.p, h->n_envs * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    QP_CUDA(cudaStreamSynchronize(h->stream));
    return QP_OK;
}

int qp_env_get_state(qp_handle* h, double* budget, int32_t* shares, double* share_value) {
    if (!h) return fail(QP_EINVAL, "null handle");
    if (!h->has_env) return fail(QP_ESTATE, "no episodes: call qp_env_create first");
    QP_TRY(use_device(h));
    if (budget) QP_CUDA(cudaMemcpyAsync(budget, h->budget.p, h->n_envs * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (shares) QP_CUDA(cudaMemcpyAsync(shares, h->shares.p, h->n_envs * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    if (share_value) QP_CUDA(cudaMemcpyAsync(share_value, h->share_value.p, h->n_envs * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    QP_CUDA(cudaStreamSynchronize(h->stream));
    return QP_OK;
}

int qp_env_set_state(qp_handle* h, const double* budget, const int32_t* shares,
                     const double* share_value, int64_t step) {
    if (!h) return fail(QP_EINVAL, "null handle");
    if (!h->has_env) return fail(QP_ESTATE, "no episodes: call qp_env_create first");
    if (step < 0 || step > h->series_len - h->hp.wn) return fail(QP_EINVAL, "step out of range");
    QP_TRY(use_device(h));
    if (budget) QP_CUDA(cudaMemcpyAsync(h->budget.p, budget, h->n_envs * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (shares) QP_CUDA(cudaMemcpyAsync(h->shares.p, shares, h->n_envs * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    if (share_value) QP_CUDA(cudaMemcpyAsync(h->share_value.p, share_value, h->n_envs * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    const int32_t s32 = (int32_t)step;
    QP_CUDA(cudaMemcpyAsync(h->d_step.p, &s32, sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    QP_CUDA(cudaStreamSynchronize(h->stream));
    h->steps_done = step;
    return QP_OK;
}

int qp_env_portfolio_stats(qp_handle* h, double* mean, double* stddev) {
    if (!h || !mean || !stddev) return fail(QP_EINVAL, "null argument");
    if (!h->has_env) return fail(QP_ESTATE, "no episodes: call qp_env_create first");
    QP_TRY(use_device(h));
    QP_TRY(compute_portfolio(h));
    const int blocks = (int)std::min<int64_t>(1024, (h->n_envs + 255) / 256);
    double host[1024];
    auto pass = [&](double shift, int pw, double* total) -> int {
        moment_kernel<<<blocks, 256, 0, h->stream>>>(h->portfolio.p, h->n_envs, shift, pw, h->blocksum.p);
        h->launches++;
        QP_CUDA(cudaGetLastError());
        QP_CUDA(cudaMemcpyAsync(host, h->blocksum.p, blocks * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        QP_CUDA(cudaStreamSynchronize(h->stream));
        double s = 0.0;
        for (int i = 0; i < blocks; ++i) s += host[i];
        *total = s;
        return QP_OK;
    };
    double sum = 0.0, var = 0.0;
    QP_TRY(pass(0.0, 1, &sum));
    const double m = sum / (double)h->n_envs;   // mean (TrainerRouterActor.scala:149)
    QP_TRY(pass(m, 2, &var));
    *mean = m;
    *stddev = std::sqrt(var / (double)h->n_envs);  // population variance (:150-151)
    return QP_OK;
}

// ---- batching front-end -------------------------------------------------------------------------
namespace {

// grows a queue to at least `need` messages (exclusive lock held: no writer is inside)
int front_grow(qp_handle* h, qp_handle::FrontQueue& q, bool update_kind, int64_t need) {
    // writers that found the queue full have pushed count past cap; every slot below cap is completely written
    if (q.count.load() > q.cap) q.count.store(q.cap);
    if (need <= q.cap) return QP_OK;
    QP_TRY(use_device(h));
    const int64_t D = h->hp.D, cap = std::max<int64_t>(std::max<int64_t>(1024, 2 * q.cap), need);
    float *st = nullptr, *nx = nullptr, *sc = nullptr;
    int64_t* ids = nullptr;
    QP_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&st), (size_t)cap * D * sizeof(float), cudaHostAllocDefault));
    QP_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&sc), (size_t)cap * sizeof(float), cudaHostAllocDefault));
    if (update_kind) QP_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&nx), (size_t)cap * D * sizeof(float), cudaHostAllocDefault));
    else QP_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&ids), (size_t)cap * sizeof(int64_t), cudaHostAllocDefault));
    const int64_t n = std::min<int64_t>(q.count.load(), q.cap);
    if (n > 0) {
        std::memcpy(st, q.states, (size_t)n * D * sizeof(float));
        std::memcpy(sc, q.scalars, (size_t)n * sizeof(float));
        if (update_kind) std::memcpy(nx, q.next, (size_t)n * D * sizeof(float));
        else std::memcpy(ids, q.ids, (size_t)n * sizeof(int64_t));
    }
    if (q.states) cudaFreeHost(q.states);
    if (q.scalars) cudaFreeHost(q.scalars);
    if (q.next) cudaFreeHost(q.next);
    if (q.ids) cudaFreeHost(q.ids);
    q.states = st; q.next = nx; q.scalars = sc; q.ids = ids; q.cap = cap;
    return QP_OK;
}

}  // namespace

int qp_enqueue_select(qp_handle* h, const float* state, int64_t state_len, float step, int64_t env_id,
                      int64_t* ticket) {
    if (!h || !state || !ticket) return fail(QP_EINVAL, "null argument");
    if (state_len != h->hp.D)
        return fail(QP_EINVAL, "SelectionAction: tensorflow input size(" + std::to_string(h->hp.D) +
                                   ") and state(" + std::to_string(state_len) + ") size do not match");
    const size_t D = (size_t)h->hp.D;
    for (;;) {
        {
            std::shared_lock<std::shared_mutex> lock(h->qrw);
            qp_handle::FrontQueue& q = h->q_sel[h->cur_sel];
            const int64_t slot = q.count.fetch_add(1);
            if (slot < q.cap) {
                std::memcpy(q.states + (size_t)slot * D, state, D * sizeof(float));
                q.scalars[slot] = step;
                q.ids[slot] = env_id;
                *ticket = h->sel_first_ticket + slot;
                return QP_OK;
            }
        }
        std::unique_lock<std::shared_mutex> lock(h->qrw);     // full: grow (the first to get here does it), then retry
        qp_handle::FrontQueue& q = h->q_sel[h->cur_sel];
        QP_TRY(front_grow(h, q, false, std::min<int64_t>(q.count.load(), q.cap) + 1));
    }
}

int qp_enqueue_update(qp_handle* h, const float* state, int64_t state_len, float reward,
                      const float* next_state, int64_t next_state_len) {
    if (!h || !state || !next_state) return fail(QP_EINVAL, "null argument");
    if (state_len != h->hp.D || next_state_len != h->hp.D)
        return fail(QP_EINVAL, "update q: tensorflow input size(" + std::to_string(h->hp.D) +
                                   ") and state(or nextState) size(" + std::to_string(state_len) + ", " +
                                   std::to_string(next_state_len) + ") do not match");
    const size_t D = (size_t)h->hp.D;
    for (;;) {
        {
            std::shared_lock<std::shared_mutex> lock(h->qrw);
            qp_handle::FrontQueue& q = h->q_upd[h->cur_upd];
            const int64_t slot = q.count.fetch_add(1);
            if (slot < q.cap) {
                std::memcpy(q.states + (size_t)slot * D, state, D * sizeof(float));
                std::memcpy(q.next + (size_t)slot * D, next_state, D * sizeof(float));
                q.scalars[slot] = reward;
                return QP_OK;
            }
        }
        std::unique_lock<std::shared_mutex> lock(h->qrw);
        qp_handle::FrontQueue& q = h->q_upd[h->cur_upd];
        QP_TRY(front_grow(h, q, true, std::min<int64_t>(q.count.load(), q.cap) + 1));
    }
}

int qp_flush(qp_handle* h, int64_t* n_selects, int64_t* n_updates) {
    if (!h) return fail(QP_EINVAL, "null handle");
    std::lock_guard<std::mutex> one_flush(h->flush_mu);
    // hand the current generations to this flush and let the actors continue into the other pair
    qp_handle::FrontQueue *qs, *qu;
    int64_t ns, nu, first_ticket;
    {
        std::unique_lock<std::shared_mutex> lock(h->qrw);
        qs = &h->q_sel[h->cur_sel]; qu = &h->q_upd[h->cur_upd];
        ns = std::min<int64_t>(qs->count.load(), qs->cap);
        nu = std::min<int64_t>(qu->count.load(), qu->cap);
        qs->count.store(ns); qu->count.store(nu);
        first_ticket = h->sel_first_ticket;
        h->sel_first_ticket += ns;
        h->cur_sel ^= 1; h->cur_upd ^= 1;
    }
    int st = QP_OK;
    if (ns > 0) {   // every queued SelectionAction is answered with the weights as they are now ...
        std::vector<int32_t> res((size_t)ns, 0);
        st = qp_select_action(h, qs->states, h->hp.D, qs->scalars, qs->ids, res.data(), nullptr, ns);
        if (st == QP_OK) {
            std::unique_lock<std::shared_mutex> lock(h->qrw);
            h->q_results.swap(res);
            h->ticket_base = first_ticket;
        }
    }
    if (st == QP_OK && nu > 0)     // ... then every queued UpdateQ becomes one gradient-SUM step
        st = qp_update_q(h, qu->states, h->hp.D, qu->scalars, qu->next, h->hp.D, nu);
    qs->count.store(0); qu->count.store(0);      // (nobody writes to the generation being flushed)
    if (st != QP_OK) return st;
    if (n_selects) *n_selects = ns;
    if (n_updates) *n_updates = nu;
    return QP_OK;
}

int qp_take_action(qp_handle* h, int64_t ticket, int32_t* action) {
    if (!h || !action) return fail(QP_EINVAL, "null argument");
    std::shared_lock<std::shared_mutex> lock(h->qrw);
    if (ticket >= h->sel_first_ticket)
        return fail(QP_ESTATE, "ticket " + std::to_string(ticket) + " is still queued: call qp_flush first");
    const int64_t idx = ticket - h->ticket_base;
    if (idx < 0 || idx >= (int64_t)h->q_results.size())
        return fail(QP_EINVAL, "ticket " + std::to_string(ticket) + " is not from the most recent flush");
    *action = h->q_results[(size_t)idx];
    return QP_OK;
}

int64_t qp_pending(qp_handle* h) {
    if (!h) return 0;
    std::shared_lock<std::shared_mutex> lock(h->qrw);
    const qp_handle::FrontQueue &qa = h->q_sel[h->cur_sel], &qb = h->q_upd[h->cur_upd];
    return std::min<int64_t>(qa.count.load(), qa.cap) + std::min<int64_t>(qb.count.load(), qb.cap);
}

// ---- checkpoint (fills the reference's saveSnapshot stub, QDecisionPolicyActor.scala:74, :91-93) -----
namespace {
struct CkptHeader {
    char magic[8];            // "QPCKPT1\0"
    int32_t window, hidden, bookkeeping, has_env;
    float epsilon, gamma, lr, bias, step_div, adagrad_init_acc;
    uint64_t seed;
    int64_t P, iteration, steps_done, n_envs, env_id_base;
    double B0;
    int32_t N0, reserved;
};
}  // namespace

int qp_save_checkpoint(qp_handle* h, const char* path) {
    if (!h || !path) return fail(QP_EINVAL, "null argument");
    QP_TRY(use_device(h));
    CkptHeader hd{};
    std::memcpy(hd.magic, "QPCKPT1", 8);
    hd.window = h->cfg.window; hd.hidden = h->cfg.hidden; hd.bookkeeping = h->cfg.bookkeeping;
    hd.has_env = h->has_env ? 1 : 0;
    hd.epsilon = h->cfg.epsilon; hd.gamma = h->cfg.gamma; hd.lr = h->cfg.lr; hd.bias = h->cfg.bias;
    hd.step_div = h->cfg.step_div; hd.adagrad_init_acc = h->cfg.adagrad_init_acc; hd.seed = h->cfg.seed;
    hd.P = h->P; hd.iteration = h->iteration; hd.steps_done = h->steps_done;
    hd.n_envs = h->has_env ? h->n_envs : 0; hd.env_id_base = h->env_id_base; hd.B0 = h->B0; hd.N0 = h->N0;
    std::vector<float> w(h->P), a(h->P);
    QP_TRY(qp_get_weights(h, w.data(), a.data()));
    std::vector<double> b, v;
    std::vector<int32_t> sh;
    if (h->has_env) {
        b.resize(h->n_envs); v.resize(h->n_envs); sh.resize(h->n_envs);
        QP_TRY(qp_env_get_state(h, b.data(), sh.data(), v.data()));
    }
    const std::string tmp = std::string(path) + ".tmp";
    FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) return fail(QP_EIO, "cannot open " + tmp + ": " + std::strerror(errno));
    bool ok = std::fwrite(&hd, sizeof hd, 1, f) == 1 && std::fwrite(w.data(), 4, w.size(), f) == w.size() &&
              std::fwrite(a.data(), 4, a.size(), f) == a.size();
    if (ok && h->has_env)
        ok = std::fwrite(b.data(), 8, b.size(), f) == b.size() && std::fwrite(sh.data(), 4, sh.size(), f) == sh.size() &&
             std::fwrite(v.data(), 8, v.size(), f) == v.size();
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) { std::remove(tmp.c_str()); return fail(QP_EIO, "short write to " + tmp); }
    if (std::rename(tmp.c_str(), path) != 0) return fail(QP_EIO, std::string("rename failed: ") + std::strerror(errno));
    return QP_OK;
}

int qp_load_checkpoint(qp_handle* h, const char* path) {
    if (!h || !path) return fail(QP_EINVAL, "null argument");
    FILE* f = std::fopen(path, "rb");
    if (!f) return fail(QP_EIO, std::string("cannot open ") + path + ": " + std::strerror(errno));
    CkptHeader hd{};
    auto bad = [&](const std::string& m) { std::fclose(f); return fail(QP_EIO, m); };
    if (std::fread(&hd, sizeof hd, 1, f) != 1 || std::memcmp(hd.magic, "QPCKPT1", 8) != 0) return bad("not a qpolicy checkpoint");
    if (hd.window != h->cfg.window || hd.hidden != h->cfg.hidden || hd.P != h->P)
        return bad("checkpoint shape (window " + std::to_string(hd.window) + ", hidden " + std::to_string(hd.hidden) +
                   ") does not match the policy");
    std::vector<float> w(h->P), a(h->P);
    if (std::fread(w.data(), 4, w.size(), f) != w.size() || std::fread(a.data(), 4, a.size(), f) != a.size())
        return bad("truncated checkpoint");
    std::vector<double> b, v;
    std::vector<int32_t> sh;
    if (hd.has_env) {
        if (!h->has_env || hd.n_envs != h->n_envs) return bad("checkpoint holds " + std::to_string(hd.n_envs) +
                                                            " episodes; create the same episodes before loading");
        b.resize(hd.n_envs); v.resize(hd.n_envs); sh.resize(hd.n_envs);
        if (std::fread(b.data(), 8, b.size(), f) != b.size() || std::fread(sh.data(), 4, sh.size(), f) != sh.size() ||
            std::fread(v.data(), 8, v.size(), f) != v.size())
            return bad("truncated checkpoint (episodes)");
    }
    std::fclose(f);
    QP_TRY(qp_set_weights(h, w.data(), a.data()));
    h->iteration = hd.iteration;
    if (hd.has_env) QP_TRY(qp_env_set_state(h, b.data(), sh.data(), v.data(), hd.steps_done));
    return QP_OK;
}

int qp_set_snapshot(qp_handle* h, const char* path, int64_t every) {
    if (!h) return fail(QP_EINVAL, "null handle");
    h->snapshot_path = path ? path : "";
    h->snapshot_every = every;
    return QP_OK;
}

int64_t qp_iteration(const qp_handle* h) { return h ? h->iteration : 0; }

// ---- price ingest (SharePriceGetter.scala:89-101, :104-108; TrainerChildActor.scala:66) ---------------
namespace {
bool parse_date_key(const std::string& s, int32_t* key) {
    // DateTimeFormatter.ofPattern("yyyy-MM-dd"), default SMART resolver: 4-digit year, 2-digit month
    // and day; a day of 29..31 beyond the month's length resolves to the month's last day.
    if (s.size() != 10 || s[4] != '-' || s[7] != '-') return false;
    for (int i : {0, 1, 2, 3, 5, 6, 8, 9}) if (s[i] < '0' || s[i] > '9') return false;
    const int y = (s[0] - '0') * 1000 + (s[1] - '0') * 100 + (s[2] - '0') * 10 + (s[3] - '0');
    const int m = (s[5] - '0') * 10 + (s[6] - '0');
    int d = (s[8] - '0') * 10 + (s[9] - '0');
    if (y < 1 || m < 1 || m > 12 || d < 1 || d > 31) return false;
    static const int len[12] = {31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31};
    int last = len[m - 1];
    if (m == 2 && ((y % 4 == 0 && y % 100 != 0) || y % 400 == 0)) last = 29;
    if (d > last) d = last;
    *key = y * 10000 + m * 100 + d;
    return true;
}

bool parse_double(const std::string& s, double* out) {
    // String.toDouble = java.lang.Double.parseDouble: optional sign, decimal / exponent forms,
    // "NaN", "Infinity", and an optional trailing f/F/d/D type suffix; hex floats are accepted too.
    if (s.empty()) return false;
    std::string t = s;
    const char lastc = t.back();
    if ((lastc == 'f' || lastc == 'F' || lastc == 'd' || lastc == 'D') && t.size() > 1 &&
        !(t.size() >= 2 && (t[0] == '0' && (t[1] == 'x' || t[1] == 'X'))))
        t.pop_back();
    if (t == "NaN" || t == "+NaN" || t == "-NaN") { *out = std::nan(""); return true; }
    if (t == "Infinity" || t == "+Infinity") { *out = INFINITY; return true; }
    if (t == "-Infinity") { *out = -INFINITY; return true; }
    for (char c : t) if (!((c >= '0' && c <= '9') || c == '.' || c == 'e' || c == 'E' || c == '+' || c == '-')) return false;
    char* end = nullptr;
    errno = 0;
    const double v = std::strtod(t.c_str(), &end);
    if (end == t.c_str() || *end != '\0') return false;
    *out = v;
    return true;
}

std::string trim_java(const std::string& s) {   // String.trim(): strips chars <= ' '
    size_t a = 0, b = s.size();
    while (a < b && (unsigned char)s[a] <= ' ') ++a;
    while (b > a && (unsigned char)s[b - 1] <= ' ') --b;
    return s.substr(a, b - a);
}
}  // namespace

int qp_parse_prices(const char* text, size_t len, float* out, int64_t capacity, int64_t* n_out) {
    if (!text || !n_out) return fail(QP_EINVAL, "null argument");
    std::map<int32_t, double> by_date;      // HashMap(parsed: _*): a repeated date keeps the LAST row
    size_t pos = 0;
    while (pos < len) {
        size_t eol = pos;
        while (eol < len && text[eol] != '\n') ++eol;
        size_t end = eol;
        if (end > pos && text[end - 1] == '\r') --end;     // getLines strips \r\n as well
        std::vector<std::string> fields;
        size_t fs = pos;
        for (size_t i = pos; i <= end; ++i)
            if (i == end || text[i] == ',') { fields.push_back(trim_java(std::string(text + fs, i - fs))); fs = i + 1; }
        while (!fields.empty() && fields.back().empty()) fields.pop_back();   // String.split drops trailing empties
        if (fields.size() >= 2) {           // case Seq(price, date) | Seq(price, date, _*)
            int32_t key; double price;
            if (parse_date_key(fields[1], &key) && parse_double(fields[0], &price)) by_date[key] = price;
        }
        pos = eol + 1;
    }
    *n_out = (int64_t)by_date.size();
    if (out) {
        if (capacity < *n_out) return fail(QP_EINVAL, "output holds " + std::to_string(capacity) + " prices, " +
                                                          std::to_string(*n_out) + " parsed");
        int64_t i = 0;
        for (const auto& kv : by_date) out[i++] = (float)kv.second;   // sorted by date; .toFloat (TrainerChildActor.scala:66)
    }
    return QP_OK;
}

// Test hook for the tensor-core GEMM alone: C[M,N] = A[M,K] · B[N,K]^T with host buffers.
int qp_debug_tc_gemm(int32_t device, int64_t M, int64_t N, int64_t K, const float* A, const float* B, float* C,
                     int32_t passes, int32_t splits, float* ms_out) {
    if (!A || !B || !C) return fail(QP_EINVAL, "null argument");
    if (M < 1 || N < 1 || K < 32 || splits < 1 || K % (32 * splits) != 0 || (passes != 1 && passes != 3))
        return fail(QP_EINVAL, "need K a multiple of 32*splits and passes in {1, 3}");
    QP_CUDA(cudaSetDevice(device));
    DevBuf<float> dA, dAh, dAl, dB, dBh, dBl, dC;
    const int tiles_m = (int)((M + kTcBM - 1) / kTcBM), tiles_n = (int)((N + kTcBN - 1) / kTcBN);
    QP_TRY(dA.ensure(M * K)); QP_TRY(dAh.ensure(M * K)); QP_TRY(dAl.ensure(M * K));
    QP_TRY(dB.ensure(N * K)); QP_TRY(dBh.ensure(N * K)); QP_TRY(dBl.ensure(N * K));
    QP_TRY(dC.ensure((size_t)splits * M * N));
    auto rel = [&]() { dA.release(); dAh.release(); dAl.release(); dB.release(); dBh.release(); dBl.release(); dC.release(); };
    int st = QP_OK;
    auto run = [&]() -> int {
        QP_CUDA(cudaMemcpy(dA.p, A, M * K * sizeof(float), cudaMemcpyHostToDevice));
        QP_CUDA(cudaMemcpy(dB.p, B, N * K * sizeof(float), cudaMemcpyHostToDevice));
        tc_split_kernel<<<(unsigned)((M * K + 255) / 256), 256>>>(dA.p, dAh.p, dAl.p, M * K);
        tc_split_kernel<<<(unsigned)((N * K + 255) / 256), 256>>>(dB.p, dBh.p, dBl.p, N * K);
        QP_CUDA(cudaGetLastError());
        CUtensorMap ta_hi, ta_lo, tb_hi, tb_lo;
        QP_TRY(make_tmap(&ta_hi, dAh.p, K, M, K * sizeof(float)));
        QP_TRY(make_tmap(&ta_lo, dAl.p, K, M, K * sizeof(float)));
        QP_TRY(make_tmap(&tb_hi, dBh.p, K, N, K * sizeof(float)));
        QP_TRY(make_tmap(&tb_lo, dBl.p, K, N, K * sizeof(float)));
        TcGemmArgs g{};
        g.nkb = (int)(K / 32 / splits); g.passes = passes;
        g.C = dC.p; g.ldc = N; g.c_split_stride = M * N; g.m_valid = (int)M; g.n_valid = (int)N;
        cudaEvent_t e0, e1;
        QP_CUDA(cudaEventCreate(&e0)); QP_CUDA(cudaEventCreate(&e1));
        QP_TRY(launch_tc_gemm(nullptr, ta_hi, ta_lo, tb_hi, tb_lo, g, tiles_m, tiles_n, splits));   // warm-up
        QP_CUDA(cudaEventRecord(e0));
        QP_TRY(launch_tc_gemm(nullptr, ta_hi, ta_lo, tb_hi, tb_lo, g, tiles_m, tiles_n, splits));
        QP_CUDA(cudaEventRecord(e1));
        QP_CUDA(cudaDeviceSynchronize());
        float ms = 0.f;
        QP_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (ms_out) *ms_out = ms;
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        std::vector<float> part((size_t)splits * M * N);
        QP_CUDA(cudaMemcpy(part.data(), dC.p, part.size() * sizeof(float), cudaMemcpyDeviceToHost));
        for (int64_t i = 0; i < M * N; ++i) {
            float s = 0.f;
            for (int z = 0; z < splits; ++z) s += part[(size_t)z * M * N + i];
            C[i] = s;
        }
        return QP_OK;
    };
    st = run();
    rel();
    return st;
}

// Test hook: single tcgen05.mma instructions on operand images a test lays out byte by byte (qp_probe.cuh).
int qp_debug_umma(int32_t device, const void* a_img, int64_t a_bytes, const void* b_img, int64_t b_bytes,
                  uint64_t desc_a, uint64_t desc_b, uint32_t idesc, int32_t n_mma, uint32_t a_step, uint32_t b_step,
                  uint32_t d_col, uint32_t d_lane, int32_t n_cols, float fill, float* out, int32_t a_tmem_k, uint32_t a_col) {
    if (!a_img || !b_img || !out) return fail(QP_EINVAL, "null argument");
    if (a_tmem_k != 0 && (a_tmem_k % 8 != 0 || a_bytes != (int64_t)128 * a_tmem_k * 4 || a_col + (uint32_t)a_tmem_k > 512))
        return fail(QP_EINVAL, "A from tensor memory: a_img must be [128][a_tmem_k] fp32 with a_tmem_k a multiple of 8");
    if (a_bytes < 16 || b_bytes < 16 || a_bytes % 16 || b_bytes % 16 || n_cols < 16 || n_cols > 512 || n_cols % 16 || n_mma < 1 ||
        d_col + (uint32_t)0 > 512)
        return fail(QP_EINVAL, "image sizes must be multiples of 16 bytes, n_cols a multiple of 16 in [16, 512]");
    const size_t smem = (((size_t)a_bytes + 1023) & ~(size_t)1023) + (size_t)b_bytes;
    QP_CUDA(cudaSetDevice(device));
    int max_smem = 0;
    QP_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    if (smem + 1024 > (size_t)max_smem) return fail(QP_EINVAL, "operand images do not fit in shared memory");
    QP_CUDA(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DevBuf<uint8_t> da, db;
    DevBuf<float> dout;
    QP_TRY(da.ensure(a_bytes)); QP_TRY(db.ensure(b_bytes)); QP_TRY(dout.ensure((size_t)128 * n_cols));
    auto rel = [&]() { da.release(); db.release(); dout.release(); };
    auto run = [&]() -> int {
        QP_CUDA(cudaMemcpy(da.p, a_img, a_bytes, cudaMemcpyHostToDevice));
        QP_CUDA(cudaMemcpy(db.p, b_img, b_bytes, cudaMemcpyHostToDevice));
        UmmaProbeArgs g{};
        g.a_img = da.p; g.a_bytes = (uint32_t)a_bytes; g.b_img = db.p; g.b_bytes = (uint32_t)b_bytes;
        g.desc_a = desc_a; g.desc_b = desc_b; g.idesc = idesc; g.n_mma = n_mma; g.a_step = a_step; g.b_step = b_step;
        g.d_col = d_col; g.d_lane = d_lane; g.n_cols = n_cols; g.fill = fill; g.out = dout.p;
        g.a_tmem = a_tmem_k != 0 ? 1 : 0; g.a_k = a_tmem_k; g.a_col = a_col;
        umma_probe_kernel<<<1, 128, smem>>>(g);
        QP_CUDA(cudaGetLastError());
        QP_CUDA(cudaDeviceSynchronize());
        QP_CUDA(cudaMemcpy(out, dout.p, (size_t)128 * n_cols * sizeof(float), cudaMemcpyDeviceToHost));
        return QP_OK;
    };
    const int st = run();
    rel();
    return st;
}

int qp_debug_phase_clocks(qp_handle* h, int32_t n_steps, int64_t* out) {
    if (!h || !out) return fail(QP_EINVAL, "null argument");
    if (!h->dbg.p || (size_t)n_steps * qp::kDbgSlots > h->dbg.cap) return fail(QP_ESTATE, "no phase stamps recorded");
    if (h->want_fine) {      // [n_steps][64]
        QP_TRY(use_device(h));
        QP_CUDA(cudaMemcpyAsync(out, h->dbg.p, (size_t)n_steps * 64 * sizeof(long long), cudaMemcpyDeviceToHost, h->stream));
        QP_CUDA(cudaStreamSynchronize(h->stream));
        return QP_OK;
    }
    QP_TRY(use_device(h));
    QP_CUDA(cudaMemcpyAsync(out, h->dbg.p, (size_t)n_steps * qp::kDbgSlots * sizeof(long long), cudaMemcpyDeviceToHost, h->stream));
    QP_CUDA(cudaStreamSynchronize(h->stream));
    return QP_OK;
}

int qp_set_option(qp_handle* h, const char* key, int64_t value) {
    if (!h || !key) return fail(QP_EINVAL, "null argument");
    if (std::strcmp(key, "phase_clocks") == 0) { h->want_dbg = value != 0; h->want_cta_ns = value == 2; h->want_fine = value == 3; return QP_OK; }
    if (std::strcmp(key, "fused") == 0) { h->allow_fused = value != 0; return QP_OK; }
    if (std::strcmp(key, "tc") == 0) { h->allow_tc = value != 0; return QP_OK; }
    if (std::strcmp(key, "p2p") == 0) { h->allow_p2p = value != 0; return QP_OK; }
    if (std::strcmp(key, "ep_tc") == 0) { h->allow_ep_tc = value != 0; return QP_OK; }
    if (std::strcmp(key, "ep_c2") == 0) { h->allow_ep_c2 = value != 0; return QP_OK; }
    if (std::strcmp(key, "ep_c2_waves") == 0) { h->allow_c2_waves = value != 0; return QP_OK; }
    if (std::strcmp(key, "fanout") == 0) { h->allow_fanout = (int)value; return QP_OK; }
    if (std::strcmp(key, "nvls") == 0) { h->allow_nvls = value != 0; return QP_OK; }
    if (std::strcmp(key, "rsag") == 0) { h->rsag_mode = (int)value; return QP_OK; }
    if (std::strcmp(key, "c2_dbg") == 0) { h->c2_dbg = (int)value; return QP_OK; }
    if (std::strcmp(key, "tc_passes") == 0) {
        if (value != 1 && value != 3) return fail(QP_EINVAL, "tc_passes must be 3 (3xTF32) or 1 (TF32)");
        h->tc_passes = (int)value;
        return QP_OK;
    }
    if (std::strcmp(key, "wide") == 0) {
        if (value == 0 && !h->tile_ok) return fail(QP_EINVAL, "this shape only runs on the GEMM path");
        if (value != 0 && (h->hp.H % 128 != 0 || h->hp.H / 32 > kWideMaxColsPerLane))
            return fail(QP_EINVAL, "the GEMM path needs hidden to be a multiple of 128 and <= 2048");
        h->wide = value != 0;
        return QP_OK;
    }
    return fail(QP_EINVAL, std::string("unknown option: ") + key);
}

int qp_get_option(qp_handle* h, const char* key, int64_t* value) {
    if (!h || !key || !value) return fail(QP_EINVAL, "null argument");
    if (std::strcmp(key, "fused") == 0) { *value = h->allow_fused ? 1 : 0; return QP_OK; }
    if (std::strcmp(key, "wide") == 0) { *value = h->wide ? 1 : 0; return QP_OK; }
    if (std::strcmp(key, "ep_tc_active") == 0) {
        *value = (h->has_env && !use_fanout(h) && use_fused(h) && (ep_tc_usable(h) || ep_c2_usable(h))) ? 1 : 0;
        return QP_OK;
    }
    if (std::strcmp(key, "ep_c2_split_active") == 0) {
        *value = (h->has_env && !use_fanout(h) && use_fused(h) && ep_c2_usable(h) && !h->prices_exact) ? 1 : 0;
        return QP_OK;
    }
    if (std::strcmp(key, "fanout_active") == 0) { *value = use_fanout(h) ? 1 : 0; return QP_OK; }
    if (std::strcmp(key, "ep_c2_active") == 0) { *value = (h->has_env && !use_fanout(h) && use_fused(h) && ep_c2_usable(h)) ? 1 : 0; return QP_OK; }
    if (std::strcmp(key, "nvls_active") == 0) {
        *value = (h->has_env && !use_fanout(h) && use_fused(h) && ep_c2_usable(h) && c2_nvls(h)) ? 1 : 0;
        return QP_OK;
    }
    if (std::strcmp(key, "rsag_active") == 0) { *value = (h->has_env && !use_fanout(h) && use_fused(h) && ep_c2_usable(h) && c2_rsag(h)) ? 1 : 0; return QP_OK; }
    if (std::strcmp(key, "p2p_active") == 0) { *value = (h->has_env && !use_fanout(h) && h->world > 1 && use_fused(h)) ? 1 : 0; return QP_OK; }
    if (std::strcmp(key, "tc_a_lo_zero") == 0) { *value = h->tc_a_lo_zero ? 1 : 0; return QP_OK; }
    if (std::strcmp(key, "tc_active") == 0) { *value = (h->wide && h->allow_tc && h->tc_ready) ? 1 : 0; return QP_OK; }
    if (std::strcmp(key, "fused_active") == 0) { *value = (h->has_env && !use_fanout(h) && use_fused(h)) ? 1 : 0; return QP_OK; }
    if (std::strcmp(key, "fused_step_active") == 0) { *value = (h->has_env && !use_fanout(h) && use_fused_step(h)) ? 1 : 0; return QP_OK; }
    return fail(QP_EINVAL, std::string("unknown option: ") + key);
}

int qp_dist_unique_id(void* out, size_t out_bytes) {
    if (!out || out_bytes < QP_DIST_ID_BYTES) return fail(QP_EINVAL, "need a 128-byte buffer");
    QP_TRY(nccl_load());
    ncclUniqueId id;
    QP_NCCL(g_nccl.GetUniqueId(&id));
    std::memcpy(out, id.internal, QP_DIST_ID_BYTES);
    return QP_OK;
}

int qp_dist_init(qp_handle* h, const void* unique_id, int32_t rank, int32_t world) {
    if (!h || !unique_id) return fail(QP_EINVAL, "null argument");
    if (world < 1 || rank < 0 || rank >= world) return fail(QP_EINVAL, "rank/world out of range");
    if (h->comm) return fail(QP_ESTATE, "qp_dist_init called twice");
    if (world == 1) { h->rank = 0; h->world = 1; return QP_OK; }
    QP_TRY(nccl_load());
    QP_TRY(use_device(h));
    ncclUniqueId id;
    std::memcpy(id.internal, unique_id, QP_DIST_ID_BYTES);
    QP_NCCL(g_nccl.CommInitRank(&h->comm, world, id, rank));
    h->rank = rank; h->world = world;
    return QP_OK;
}

int qp_dist_p2p_export(qp_handle* h, void* out, size_t out_bytes) {
    if (!h || !out || out_bytes < QP_DIST_IPC_BYTES) return fail(QP_EINVAL, "need a 64-byte buffer");
    static_assert(sizeof(cudaIpcMemHandle_t) == QP_DIST_IPC_BYTES, "IPC handle size");
    QP_TRY(use_device(h));
    if (!h->p2p_block) {
        const size_t bytes = p2p_tc_offset(h) + (size_t)2 * 8 * p2p_pad(h) * sizeof(unsigned long long);
        QP_CUDA(cudaMalloc(&h->p2p_block, bytes));
        QP_CUDA(cudaMemset(h->p2p_block, 0, bytes));
        QP_CUDA(cudaDeviceSynchronize());
    }
    cudaIpcMemHandle_t hd;
    QP_CUDA(cudaIpcGetMemHandle(&hd, h->p2p_block));
    std::memcpy(out, &hd, sizeof hd);
    return QP_OK;
}

int qp_dist_p2p_import(qp_handle* h, const void* handles, int32_t world) {
    if (!h || !handles) return fail(QP_EINVAL, "null argument");
    if (world != h->world || world > 8) return fail(QP_EINVAL, "world must match qp_dist_init and be <= 8");
    if (!h->p2p_block) return fail(QP_ESTATE, "call qp_dist_p2p_export first");
    QP_TRY(use_device(h));
    for (int r = 0; r < world; ++r) {
        if (r == h->rank) { h->p2p_peer_block[r] = h->p2p_block; continue; }
        cudaIpcMemHandle_t hd;
        std::memcpy(&hd, static_cast<const char*>(handles) + (size_t)r * QP_DIST_IPC_BYTES, sizeof hd);
        void* ptr = nullptr;
        QP_CUDA(cudaIpcOpenMemHandle(&ptr, hd, cudaIpcMemLazyEnablePeerAccess));
        h->p2p_peer_block[r] = ptr;
    }
    h->p2p_ready = true;      // (p2p_flag_base keeps counting: the tags left in the blocks stay in the past)
    return QP_OK;
}

int64_t qp_dist_p2p_bytes(const qp_handle* h) {
    if (!h) return 0;
    return (int64_t)(p2p_tc_offset(h) + (size_t)2 * 8 * p2p_pad(h) * sizeof(unsigned long long));
}

int qp_dist_p2p_attach(qp_handle* h, const uint64_t* block_ptrs, int32_t world, uint64_t multicast_ptr) {
    if (!h || !block_ptrs) return fail(QP_EINVAL, "null argument");
    if (world != h->world || world > 8 || world < 2) return fail(QP_EINVAL, "world must match qp_dist_init and be 2..8");
    if (h->p2p_block) return fail(QP_ESTATE, "an exchange block exists already (qp_dist_p2p_export)");
    for (int r = 0; r < world; ++r) {
        if (!block_ptrs[r] || (block_ptrs[r] & 255u)) return fail(QP_EINVAL, "exchange blocks must be non-null and 256-byte aligned");
        h->p2p_peer_block[r] = reinterpret_cast<void*>(block_ptrs[r]);
    }
    h->p2p_block = h->p2p_peer_block[h->rank];
    h->p2p_external = true;
    h->p2p_mc = reinterpret_cast<void*>(multicast_ptr);
    h->p2p_ready = true;
    return QP_OK;
}

int qp_dist_nccl_version(int32_t* version) {
    if (!version) return fail(QP_EINVAL, "null argument");
    QP_TRY(nccl_load());
    int v = 0;
    if (g_nccl.GetVersion) QP_NCCL(g_nccl.GetVersion(&v));
    *version = v;
    return QP_OK;
}

int qp_measure_fp32_peak(qp_handle* h, double* tflops) {
    if (!h || !tflops) return fail(QP_EINVAL, "null argument");
    QP_TRY(use_device(h));
    const int blocks = h->num_sms * 8, iters = 1 << 14;
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        QP_CUDA(cudaEventRecord(h->ev[0], h->stream));
        fma_peak_kernel<<<blocks, 256, 0, h->stream>>>(h->grad.p, iters, 1.0000001f, 1e-9f);
        h->launches++;
        QP_CUDA(cudaGetLastError());
        QP_CUDA(cudaEventRecord(h->ev[1], h->stream));
        QP_CUDA(cudaEventSynchronize(h->ev[1]));
        float ms = 0.0f;
        QP_CUDA(cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]));
        const double flops = 2.0 * 16.0 * (double)iters * 256.0 * (double)blocks;
        if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    *tflops = best;
    return QP_OK;
}

int qp_measure_fp32_outer_peak(qp_handle* h, double* tflops) {
    if (!h || !tflops) return fail(QP_EINVAL, "null argument");
    QP_TRY(use_device(h));
    const int blocks = h->num_sms * 8, iters = 1 << 12;
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        QP_CUDA(cudaEventRecord(h->ev[0], h->stream));
        fma_outer_kernel<<<blocks, 256, 0, h->stream>>>(h->grad.p, iters);
        h->launches++;
        QP_CUDA(cudaGetLastError());
        QP_CUDA(cudaEventRecord(h->ev[1], h->stream));
        QP_CUDA(cudaEventSynchronize(h->ev[1]));
        float ms = 0.0f;
        QP_CUDA(cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]));
        const double flops = 2.0 * 64.0 * (double)iters * 256.0 * (double)blocks;
        if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    *tflops = best;
    return QP_OK;
}

int qp_train(qp_handle* h, const float* prices, int64_t series_len, int64_t n_series,
             const int32_t* series_of_env, int64_t n_envs, int64_t env_id_base, double budget0,
             int32_t shares0, double* portfolios_out) {
    if (!portfolios_out) return fail(QP_EINVAL, "null portfolios_out");
    QP_TRY(qp_env_create(h, prices, series_len, n_series, series_of_env, n_envs, env_id_base, budget0, shares0));
    const int64_t total = qp_env_steps_total(h);
    QP_TRY(qp_env_step_async(h, (int32_t)total));
    return qp_env_get_portfolio(h, portfolios_out);
}

}  // extern "C"

// =================================================================================================
// Allocation-free TD gradient.
//
// Who this is for: code that must not reach the CUDA allocator when it calls into this library.
// Two kinds of caller of this policy are in that position.  (1) An actor running on a shared
// dispatcher next to other GPU work: cudaMalloc/cudaFree serialise the whole device, so a scratch
// buffer that grows inside an UpdateQ stalls every other actor's stream; the handle-based entry
// points above may do exactly that the first time a batch is larger than any seen before.
// (2) A host process whose device memory has been handed over in full to another framework's pool
// (a JVM-side tensor runtime, an inference server): a cudaMalloc from here simply fails, while the
// caller can still pass us pointers that it carved out of that pool.
//
// The scratch this path needs (the hidden-layer backward signal and the per-CTA gradient partials)
// therefore lives in one file-scope device array of fixed size that is part of the library image:
// each GPU that loads libqpolicy.so carries it from load until the process exits.
// =================================================================================================
namespace {

constexpr size_t kStaticScratchFloats = size_t(1) << 20;   // 4 MiB per GPU, stated in qpolicy.h
constexpr int kStaticMaxSplit = 16;                        // dW1 partials kept by this variant
constexpr int kStaticMaxDevices = 64;

static __device__ __align__(256) float g_static_scratch[kStaticScratchFloats];

struct StaticDevice {
    bool ready = false;
    float* scratch = nullptr;   // DEVICE address of g_static_scratch on this GPU
    int num_sms = 0;
    int max_smem = 0;
};
StaticDevice g_static_dev[kStaticMaxDevices];
std::mutex g_static_mutex;

// Once per device, with that device current.  In host code the name g_static_scratch is only a
// stand-in for the symbol — handing it to a kernel would fault (or, where the GPU can reach host
// memory, quietly use the wrong bytes) — so the device address is looked up here and kept.
int static_setup(int device, StaticDevice** out) {
    std::lock_guard<std::mutex> lock(g_static_mutex);
    StaticDevice& d = g_static_dev[device];
    if (!d.ready) {
        float* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_static_scratch);
        if (e != cudaSuccess || p == nullptr)
            return fail(QP_ECUDA, std::string("cudaGetSymbolAddress(g_static_scratch): ") + cudaGetErrorString(e));
        QP_CUDA(cudaDeviceGetAttribute(&d.num_sms, cudaDevAttrMultiProcessorCount, device));
        QP_CUDA(cudaDeviceGetAttribute(&d.max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
        QP_CUDA(cudaFuncSetAttribute(step_kernel<kMsgUpdate>, cudaFuncAttributeMaxDynamicSharedMemorySize, d.max_smem));
        QP_CUDA(cudaFuncSetAttribute(dw1_kernel<kMsgUpdate>, cudaFuncAttributeMaxDynamicSharedMemorySize, d.max_smem));
        d.scratch = p;
        d.ready = true;
    }
    *out = &d;
    return QP_OK;
}

}  // namespace

extern "C" int qp_td_grad_noalloc(int32_t device, const qp_config* cfg, const float* d_params,
                                  const float* d_states, const float* d_rewards,
                                  const float* d_next_states, int64_t B, float* d_grad_out,
                                  void* stream) {
    if (!cfg || !d_params || !d_states || !d_rewards || !d_next_states || !d_grad_out)
        return fail(QP_EINVAL, "null argument");
    if (device < 0 || device >= kStaticMaxDevices) return fail(QP_EINVAL, "device ordinal out of range");
    if (cfg->window < 1 || cfg->hidden < 4 || cfg->hidden % 4)
        return fail(QP_EINVAL, "window >= 1 and hidden a multiple of 4 (>= 4) required");
    if (B < 1) return fail(QP_EINVAL, "B < 1");
    // The kernels fetch W1 with bulk copies (16-byte granules) and the rows with 4-byte loads: a pointer carved out
    // of somebody else's pool at an odd offset would fault inside the kernel, so it is refused here instead.
    if ((reinterpret_cast<uintptr_t>(d_params) & 15u) != 0)
        return fail(QP_EINVAL, "qp_td_grad_noalloc: d_params must be 16-byte aligned");
    if (((reinterpret_cast<uintptr_t>(d_states) | reinterpret_cast<uintptr_t>(d_rewards) |
          reinterpret_cast<uintptr_t>(d_next_states) | reinterpret_cast<uintptr_t>(d_grad_out)) & 3u) != 0)
        return fail(QP_EINVAL, "qp_td_grad_noalloc: device pointers must be 4-byte aligned");
    // the caller's current device is put back on every way out
    struct DeviceGuard {
        int prev = -1;
        ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    } guard;
    if (cudaGetDevice(&guard.prev) != cudaSuccess) guard.prev = -1;
    QP_CUDA(cudaSetDevice(device));
    StaticDevice* dev = nullptr;
    QP_TRY(static_setup(device, &dev));

    Hyper hp{};
    hp.wn = cfg->window; hp.D = cfg->window + 2; hp.H = cfg->hidden;
    hp.Hs = (int)round_up(cfg->hidden, kColsPerWarp);
    hp.epsilon = cfg->epsilon; hp.gamma = cfg->gamma; hp.lr = cfg->lr; hp.bias = cfg->bias;
    hp.step_div = cfg->step_div; hp.bookkeeping = cfg->bookkeeping; hp.seed = cfg->seed;
    const int64_t P1 = (int64_t)hp.D * hp.H, P = P1 + 3LL * hp.H;
    Geometry geo;
    QP_TRY(make_geometry(hp, dev->max_smem, &geo));

    // carve the fixed array: [dz1 | dW1 partials | dW2 partials], each piece 256-byte aligned
    const int64_t Bpad = round_up(B, kTileEnvs);
    const int g1 = (int)std::min<int64_t>(Bpad / kTileEnvs, dev->num_sms);
    const int g2 = std::min(dw1_splits(B, dev->num_sms, geo.dw1_tiles), kStaticMaxSplit);
    const size_t n_dz1 = (size_t)round_up(Bpad * hp.H, 64);
    const size_t n_p1 = (size_t)round_up((int64_t)g2 * P1, 64);
    const size_t n_p2 = (size_t)round_up((int64_t)g1 * 3 * hp.H, 64);
    if (n_dz1 + n_p1 + n_p2 > kStaticScratchFloats)
        return fail(QP_EINVAL, "qp_td_grad_noalloc: B = " + std::to_string(B) + " rows at window " +
                                   std::to_string(hp.wn) + ", hidden " + std::to_string(hp.H) + " need " +
                                   std::to_string((n_dz1 + n_p1 + n_p2) * sizeof(float)) +
                                   " bytes of scratch; the built-in array holds " +
                                   std::to_string(kStaticScratchFloats * sizeof(float)) +
                                   " (use a smaller batch, or qp_update_q which allocates)");
    float* dz1 = dev->scratch;
    float* part1 = dz1 + n_dz1;
    float* part2 = part1 + n_p1;
    cudaStream_t st = (cudaStream_t)stream;

    // Nothing in the array has to start at zero: every word the reduction reads is written earlier
    // in this same call (K1 writes all of dz1 and its dW2 partial, K2 writes whole D x H partials),
    // so whatever the previous call left behind is never observed.
    StepArgs s{};
    s.hp = hp; s.params = d_params;
    s.states = d_states; s.next_states = d_next_states; s.rewards = d_rewards;
    s.B = B; s.Bpad = Bpad; s.dz1 = dz1; s.dw2_part = part2;
    step_kernel<kMsgUpdate><<<g1, geo.step_warps * 32, geo.smem_step[kMsgUpdate], st>>>(s);
    QP_CUDA(cudaGetLastError());

    Dw1Args d{};
    d.hp = hp; d.states = d_states; d.Bpad = Bpad; d.dz1 = dz1; d.B = B; d.part = part1;
    d.n_rt = geo.n_rt; d.cg_max = geo.cg_max;
    dw1_kernel<kMsgUpdate><<<dim3(g2, geo.dw1_tiles), geo.dw1_threads, geo.smem_dw1, st>>>(d);
    QP_CUDA(cudaGetLastError());

    ApplyArgs a{};
    a.hp = hp; a.grad = d_grad_out;
    a.part1 = part1; a.n1 = g2; a.part2 = part2; a.n2 = g1;
    a.P = P; a.P1 = P1; a.do_reduce = 1; a.do_adagrad = 0;
    apply_kernel<<<(int)((P * kApplyLanes + 255) / 256), 256, 0, st>>>(a);
    QP_CUDA(cudaGetLastError());
    QP_CUDA(cudaStreamSynchronize(st));
    return QP_OK;
}
