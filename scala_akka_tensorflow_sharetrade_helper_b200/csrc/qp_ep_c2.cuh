// qp_ep_c2.cuh — the persistent episode kernel on thread-block clusters of two: a CTA pair owns 64 episodes and
// splits the HIDDEN dimension, CTA r of the pair taking columns [r*H/2, (r+1)*H/2).  DESIGN.md §4 (i-c2).
//
// Why: in episode_tc_kernel (qp_ep_tc.cuh) every CTA streams all of W1^T (hi + lo, 372 KB per env-step at the
// reference shape) for 32 episodes whose s / s' rows fill half of the M = 128 MMA, and produces a full-width dW1
// partial per 32 episodes.  With the hidden dimension split over a pair
//   * each SM takes in HALF of W1^T per step (the forward was bound by that per-SM ingest),
//   * the M = 128 rows of the forward are all live (64 x s + 64 x s'),
//   * the backward contracts over 64 episodes: half as many dW1 partials to store, exchange and add.
// The two CTAs need each other only for the output layer: Q(s) and Q(s') are sums over ALL hidden columns, so each
// CTA adds up its half and the pair swaps 12 partial sums per episode through distributed shared memory.
//
// Operand layouts (tcgen05, kind::tf32; facts pinned by tests/test_gpu_umma_layouts.py):
//   * the price windows live in shared memory as X^T, [256 rows = time][32 episodes] per 32-episode block, 128-byte
//     rows, SWIZZLE_128B computed from the absolute address: the K-major A operand of the BACKWARD contraction
//     (M = time, K = episodes) whose start may be ANY row, so the window slides by moving the descriptor's start
//     address one row per env-step (rows are copied back to the top every kC2Slide steps);
//   * kind::tf32 has no usable MN-major form of such a buffer (only the SWIZZLE_128B_BASE32B pattern, which no
//     K-major descriptor can read), so the FORWARD A operand X (M = episodes and s / s', K = time) is transposed out
//     of the window buffer per 32-wide k-block into a 3-stage ring by eight builder warps while the tensor core
//     consumes the previous k-block;
//   * B of the forward = this CTA's rows of W1^T (tf32 hi / lo), streamed by TMA as before; B of the backward =
//     dZ1 (hi / lo), written by the element-wise phase.
// Everything else — select / book-keeping / TD arithmetic, the gradient all-to-all (phase B), the NVLink share
// exchange, the split grid barrier — is episode_tc_kernel's, on 64-episode tiles, with three differences: the dW1 partial
// is stored transposed (c2_kk below), prices that are not tf32-exact are taken as two episodes each (kSplit), and more
// tiles than pairs run in waves whose backward sums stay in TMEM (kWaves); see qp_ep_c2_kernel.cuh.
#pragma once
#include "qp_ep_tc.cuh"
#include <type_traits>

namespace qp {

constexpr int kC2Threads = 384;
constexpr int kC2Warps = kC2Threads / 32;
constexpr int kC2Envs = 64;             // episodes per CTA pair
constexpr int kC2Rows = 256;            // rows of the window buffer
constexpr int kC2Slide = 48;            // env-steps between copy-backs (a multiple of 8: the swizzle phase of a row is kept)
constexpr int kC2AStages = 3;           // forward A operand ring: [128 rows x 32 k] tiles of 16 KB
constexpr int kC2WStages = 3;           // W1^T ring: (hi, lo) tile pairs of 2 * Nh * 128 bytes
constexpr int kC2Builders = 8;          // warps 2..9 transpose the A tiles
constexpr int kC2Sub = kC2Warps / 2;    // warps per 32-episode lane group in the element-wise phases
constexpr int kC2EStride = 65;          // padded row stride of the [column][episode] tiles

// The dW1 partial of a pair is stored TRANSPOSED, [hidden column][input row] with rows of c2_kk(D) floats (D rounded up to
// 4): phase B owns contiguous ranges of it, so consecutive threads own consecutive window rows k of one column n — which
// is the order W1^T is laid out in: the tf32 hi / lo of the new weights go out as coalesced 128-byte stores (with the
// partial in W1's own [row][column] order they were 4-byte stores 896 bytes apart: 2.6 MB of sector writes per step).
__host__ __device__ inline int c2_kk(int D) { return (D + 3) & ~3; }

struct EpC2Args {
    EpisodeArgs ea;
    float* w1T_hi; float* w1T_lo;   // [H, Kp] K-major tf32 hi/lo copies of W1[0:wn, :] (as episode_tc_kernel)
    int Kp, nkb;
    int Hh;                          // hidden columns per CTA (H / 2)
    int Nh;                          // MMA N per CTA: Hh rounded up to 8
    unsigned long long* published;
    unsigned long long* pdata[8];    // several GPUs: every rank's exchange block (see EpTcArgs)
    int rsag;                        // 1: reduce-scatter + all-gather (two NVLink crossings, 1/4 of the bytes at eight ranks), 0: all-to-all
    unsigned long long* mc_pdata;    // the same block through its multicast address (NVLS), or null: one store reaches every rank
    int64_t p_pad;
    int dbg_mode;                    // timing experiments (option "c2_dbg"); 0 in every real run
    int win_tma;                     // waves: 1 = episode e is series e and tX maps the transposed price matrix [T][series]
    int split;                       // 1: prices not tf32-exact, 32 real + 32 virtual episodes per pair (host picks the instantiation)
    long long* cta_ns;               // tuning aid ("phase_clocks" = 2): [n_steps][gridDim][8] globaltimer stamps of every CTA
    long long* fine;                 // tuning aid ("phase_clocks" = 3): [n_steps][64] clock64 stamps inside the forward of CTA 0
};

// Shared-memory map (bytes from the 1024-aligned base); the host computes the same numbers.
//   xt   : the window buffer, two 32-episode blocks of 256 rows x 128 bytes
//   op   : the operand region.  Forward: A ring (kC2AStages x 16 KB) then W1^T ring (kC2WStages x 2 TBh).  Afterwards:
//          z / h tiles at op, the twelve partial sums per warp (qpart) and then dZ1 hi / lo at op + bwd; then the
//          epilogue's staging boxes; then phase B's stage.
struct C2Smem {
    size_t xt, op, op_bytes, wring, bwd, stg_a, stg_b, w2s, wb, ws, xe, dz2s, kns, qsum, part5, bars, total, TBh;
};
__host__ __device__ inline C2Smem c2_smem(int Hh, int Nh) {
    C2Smem m{};
    m.TBh = (size_t)Nh * 128;
    m.xt = 0;
    m.op = (size_t)2 * kC2Rows * 128;
    m.wring = m.op + (size_t)kC2AStages * 16384;
    const size_t hz = (((size_t)2 * Hh * kC2EStride * 4) + 1023) & ~(size_t)1023;
    m.bwd = m.op + hz;                                          // dZ1 hi / lo (4 TBh), qpart before them
    // the epilogue's staging boxes (4 KB per warp, 1024-byte aligned): six where z1(s') was, six behind dZ1
    m.stg_a = (((size_t)Hh * kC2EStride * 4) + 1023) & ~(size_t)1023;
    m.stg_b = hz + 4 * m.TBh;
    size_t need = m.stg_b + (size_t)(kC2Warps - 6) * 4096;
    if (m.stg_a + 6 * 4096 > hz) need = (size_t)1 << 30;      // (shape outside the kernel: the host refuses it)
    const size_t fwd = (size_t)kC2AStages * 16384 + (size_t)kC2WStages * 2 * m.TBh;
    if (fwd > need) need = fwd;
    m.op_bytes = (need + 1023) & ~(size_t)1023;
    size_t o = m.op + m.op_bytes;
    m.w2s = o; o += (size_t)Hh * 16;
    m.wb = o; o += (size_t)Hh * 4;
    m.ws = o; o += (size_t)Hh * 4;
    m.xe = o; o += 2 * kC2Envs * 4;
    m.dz2s = o; o += kC2Envs * 4;
    m.kns = o; o += kC2Envs * 4;
    o = (o + 15) & ~(size_t)15;
    m.qsum = o; o += (size_t)2 * 12 * kC2Envs * 4;
    m.part5 = o; o += (size_t)2 * 5 * Hh * 4;
    o = (o + 7) & ~(size_t)7;
    m.bars = o; o += 24 * 8;
    m.total = o + 16;
    return m;
}

#ifdef __CUDACC__

__device__ __forceinline__ void c2_mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ uint32_t c2_cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void c2_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
// One of the twelve output-layer sums of an episode: this CTA's half plus the partner's, in an order both CTAs share
// (kSplit: the two column subsets of the real episode's lane groups first).
template <bool kSplit>
__device__ __forceinline__ float c2_qsum(const float* qsum, int j, int el, int lane) {
    if constexpr (kSplit)
        return (qsum[j * kC2Envs + lane] + qsum[j * kC2Envs + 32 + lane]) + (qsum[(12 + j) * kC2Envs + lane] + qsum[(12 + j) * kC2Envs + 32 + lane]);
    else
        return qsum[j * kC2Envs + el] + qsum[(12 + j) * kC2Envs + el];
}
// address of the same shared-memory variable in the CTA with the given rank of this cluster
__device__ __forceinline__ uint32_t c2_map_peer(const void* p, uint32_t rank) {
    uint32_t a = (uint32_t)__cvta_generic_to_shared(p), r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
    return r;
}
__device__ __forceinline__ void c2_st_cluster(uint32_t addr, float v) {
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void c2_tma_store_4d(const CUtensorMap* tmap, const void* smem_src, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];\n" ::"l"(tmap),
                 "r"((uint32_t)__cvta_generic_to_shared(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}

// The share exchange of phase B between the GPUs of a node (world > 1): takes this rank's sum of parameter p, returns the
// sum over all ranks (the same bits everywhere).  kRsag picks the scheme at compile time: the kernel sits at its register
// limit, and carrying both schemes (or a call) costs spills that slow every phase.
template <bool kRsag>
__device__ __forceinline__ float c2_share_exchange(const EpC2Args& x, float g, int64_t p, int32_t sidx, int cta) {
    const EpisodeArgs& ea = x.ea;
    if constexpr (kRsag) {
            // ---- gradient all-reduce as reduce-scatter + all-gather over NVLink peer memory (eight ranks: the all-to-all
            // below makes every GPU take in world - 1 copies of the whole gradient per step, 2.3 MB, which is what
            // bounds it).  The chunk of CTA c belongs to rank c mod world.  Every other rank sends its local sum there,
            // {value, step tag} in one 64-bit store; the owner adds the world values in rank order and sends the result
            // back to everybody (one multicast store where there is NVLS), tagged the same way.  No fence, no flag; two
            // NVLink crossings, 0.6 MB taken in per GPU and step.
            const int owner = cta % ea.world;
            const unsigned int want = ea.flag_base + (unsigned int)sidx + 1u;
            const size_t slot = (size_t)(want & 1u) * (size_t)ea.world;
            const unsigned long long* local = x.pdata[ea.rank] + slot * (size_t)x.p_pad + p;
            unsigned long long t0 = 0;
            auto overdue = [&](unsigned int spins) {
                if ((spins & 4095u) == 4095u) {
                    unsigned long long t;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                    if (t0 == 0) t0 = t;
                    else if (t - t0 > 20000000000ull) __trap();   // a rank never arrived: fail instead of hanging
                }
            };
            if (ea.rank != owner) {
                const unsigned long long mine = ((unsigned long long)want << 32) | (unsigned long long)__float_as_uint(g);
                asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(x.pdata[owner] + (slot + (size_t)ea.rank) * (size_t)x.p_pad + p),
                             "l"(mine) : "memory");
                for (unsigned int spins = 0;; ++spins) {        // the owner's sum arrives in the plane of the owner's rank
                    unsigned long long v;
                    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(local + (size_t)owner * (size_t)x.p_pad) : "memory");
                    if ((unsigned int)(v >> 32) == want) { g = __uint_as_float((unsigned int)v); break; }
                    overdue(spins);
                }
            } else {
                float pv[8];
#pragma unroll
                for (int r = 0; r < 8; ++r) pv[r] = 0.0f;
                pv[ea.rank & 7] = g;
                unsigned int pending = ((1u << ea.world) - 1u) & ~(1u << ea.rank);
                for (unsigned int spins = 0; pending != 0u; ++spins) {
                    unsigned long long vv[8];
#pragma unroll
                    for (int r = 0; r < 8; ++r)
                        if ((pending >> r) & 1u)
                            asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(vv[r]) : "l"(local + (size_t)r * (size_t)x.p_pad) : "memory");
#pragma unroll
                    for (int r = 0; r < 8; ++r)
                        if (((pending >> r) & 1u) && (unsigned int)(vv[r] >> 32) == want) {
                            pv[r] = __uint_as_float((unsigned int)vv[r]);
                            pending &= ~(1u << r);
                        }
                    overdue(spins);
                }
                g = 0.0f;
#pragma unroll
                for (int r = 0; r < 8; ++r) g += pv[r];
                const unsigned long long res = ((unsigned long long)want << 32) | (unsigned long long)__float_as_uint(g);
                if (x.mc_pdata) {
                    asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" ::"l"(x.mc_pdata + (slot + (size_t)owner) * (size_t)x.p_pad + p),
                                 "l"(res) : "memory");
                } else {
#pragma unroll
                    for (int r = 0; r < 8; ++r)
                        if (r < ea.world && r != ea.rank)
                            asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(x.pdata[r] + (slot + (size_t)owner) * (size_t)x.p_pad + p),
                                         "l"(res) : "memory");
                }
            }
    } else {

            // ---- gradient all-reduce over NVLink peer memory, share by share and without a fence: every value travels
            // as one 8-byte store {value, step tag} into the slot this rank owns in every peer's exchange block (posted
            // writes, one NVLink crossing); the receiver spins on its LOCAL slots until the tag is this step's, then adds
            // the values in rank order — the same bits on every rank.  Two buffers by (global) step parity.
            const unsigned int want = ea.flag_base + (unsigned int)sidx + 1u;
            const size_t slot = (size_t)(want & 1u) * (size_t)ea.world;
            const unsigned long long mine = ((unsigned long long)want << 32) | (unsigned long long)__float_as_uint(g);
            if (x.mc_pdata) {
                // NVLS: one store to the multicast address; the switch delivers it into the slot [parity][this rank][p] of
                // EVERY rank's block (this rank's included): one NVLink crossing of 8 bytes instead of world - 1
                asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" ::"l"(x.mc_pdata + (slot + (size_t)ea.rank) * (size_t)x.p_pad + p),
                             "l"(mine) : "memory");
            } else {
#pragma unroll
                for (int r = 0; r < 8; ++r)
                    if (r < ea.world && r != ea.rank)
                        asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(x.pdata[r] + (slot + (size_t)ea.rank) * (size_t)x.p_pad + p),
                                     "l"(mine) : "memory");
            }
            const unsigned long long* local = x.pdata[ea.rank] + slot * (size_t)x.p_pad + p;
            float pv[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) pv[r] = 0.0f;
            pv[ea.rank & 7] = g;
            unsigned int pending = ((1u << ea.world) - 1u) & ~(1u << ea.rank);
            unsigned long long t0 = 0;
            for (unsigned int spins = 0; pending != 0u; ++spins) {
                unsigned long long vv[8];
#pragma unroll
                for (int r = 0; r < 8; ++r)
                    if ((pending >> r) & 1u)
                        asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(vv[r])
                                     : "l"(local + (size_t)r * (size_t)x.p_pad) : "memory");
#pragma unroll
                for (int r = 0; r < 8; ++r)
                    if (((pending >> r) & 1u) && (unsigned int)(vv[r] >> 32) == want) {
                        pv[r] = __uint_as_float((unsigned int)vv[r]);
                        pending &= ~(1u << r);
                    }
                if ((spins & 4095u) == 4095u) {
                    unsigned long long t;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                    if (t0 == 0) t0 = t;
                    else if (t - t0 > 20000000000ull) __trap();   // a rank never arrived: fail instead of hanging
                }
            }
            g = 0.0f;
#pragma unroll
            for (int r = 0; r < 8; ++r) g += pv[r];
            }
    return g;
}

#endif  // __CUDACC__
}  // namespace qp

#include "qp_ep_c2_kernel.cuh"
