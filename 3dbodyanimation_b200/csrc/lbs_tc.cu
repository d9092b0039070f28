// Entry (1), float32, large batches: shape blend on the tensor cores, skinning on the CUDA cores, one body per TMEM lane.
// Replaces the per-vertex part of ark::Avatar::update (external/avatar/Avatar.cpp:26 shape blend, :69-73 skinning).
//
//   cloud[b][v] = sum_k weight[v][k] * ( R[b][j_vk] * (template[v] + shapedirs[v] w[b]) + t[b][j_vk] )
//
// CTA = 128 bodies x a run of sort tiles (64 consecutive vertices each; two MMA tiles of 32 slots).
//   * shape blend: D[128 bodies x 96 (slot, coordinate)] = A[128 x 32] B[96 x 32]^T by tcgen05.mma kind::tf32 into TMEM.
//     K = 32 carries a TF32 hi/lo split so that the result is good to ~2^-22:  k 0-9 w_hi * S_hi, 10-19 w_hi * S_lo,
//     20-29 w_lo * S_hi, 30 1 * T_hi, 31 1 * T_lo (S = shapedirs, T = template).  A lives in TMEM (columns 288-319, written once
//     per CTA); B arrives per MMA tile as a TMA bulk copy (cp.async.bulk + mbarrier) of a blob the model build laid out in the
//     canonical K-major / no-swizzle layout (internal.hpp TcTables); two accumulator buffers in TMEM (columns 320-511).
//   * a lane of a skinning warp is a body (tcgen05.ld 32x32b: thread <-> TMEM lane).  The body's 24 joint transforms live in
//     TMEM columns 0-287 (written once per CTA by tcgen05.st); the four of the current joint set sit in 48 registers and are
//     re-read only for the register sets the host marked (slots are ordered so that consecutive joint sets overlap, and the sets
//     carry over from sort tile to sort tile).  Per slot the warp reads one 32-byte record (weights, position, control word) as
//     warp-uniform LDS.128.
//   * results go to a [32 bodies][3 x 64 + 2] staging tile per TMEM lane quadrant, double buffered by sort tile; a finished tile
//     leaves as TMA bulk row stores (768 B per body) while the next one is computed: HBM sees the cloud write only.  The stores
//     are issued by two store warps (mbarrier hand-over out_full / out_free per quadrant and buffer): a bulk store holds its
//     issuing warp until the SM's store path has taken it, and a skinning warp that waits there idles its scheduler.
// Warps: 16 skinning warps (TMEM lane quadrant = warp & 3, slot sequence = warp >> 2: four warps per scheduler hide the FMA / LDS /
// TMEM latencies) + a service warpgroup (warp 16 lane 0 issues the TMA copies, warp 17 lane 0 the MMAs, warps 18 / 19 the row
// stores of quadrants 0, 2 / 1, 3) that hands its registers to the skinning warps with setmaxnreg (which only redistributes what
// the CTA's own warps release).  Measurements and the variants that lost: DESIGN.md sections 4, 6 and 8.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "internal.hpp"

namespace b3d {
namespace {

constexpr int TC_NST = 4;
constexpr int TC_WPQ = TC_SEQS;                  // epilogue warps per TMEM lane quadrant (= slot sequences per sort tile)
constexpr int TC_SPW = TC_SLOTS / TC_WPQ;        // slots of an MMA tile per warp
constexpr int TC_EPI_WARPS = 4 * TC_WPQ;
constexpr int TC_THREADS = 32 * (TC_EPI_WARPS + 4);   // + a service warpgroup: warp 16 issues the TMA copies, warp 17 the MMAs
constexpr int TC_EPI_THREADS = 32 * TC_EPI_WARPS;
constexpr int TC_M = 128;
constexpr int TC_SROW = 3 * TC_T + 2;           // 194 floats (see TcSmem::out)
constexpr int TC_COL_D0 = 320, TC_COL_D1 = 416;  // accumulator buffers; columns [0, 288) hold the joint transforms
constexpr int TC_A_SLAB = TC_M * 32, TC_B_SLAB = TC_N * 32;   // bytes per K = 8 step of the operands

constexpr int TC_GROW = NJ * 12 + 4;            // 292 floats: prologue rows of joint transforms, 16 B aligned and conflict free for LDS.128
constexpr int TC_NREC = 4;                       // ring of slot-record blocks (1 KB per MMA tile)
constexpr int TC_NOP = 2;                        // ring of B operand tiles (12 KB per MMA tile)
constexpr int TC_REC_BYTES = TC_BLOB_BYTES - TC_OPER_BYTES;
constexpr int TC_CH_A = TC_M * 16 * 4, TC_CH_B = TC_N * 16 * 4, TC_CH_BYTES = TC_CH_A + TC_CH_B;   // 8192 + 6144
constexpr int TC_NCH = 2 + TC_POSE_KSTEPS / 2;   // chunks per MMA tile with pose blendshapes: 2 shape + 13 pose
static_assert(TC_POSE_KSTEPS % 2 == 0, "pose K steps come in pairs");
constexpr int TC_COL_A = 288;                    // A operand (128 bodies x K = 32) lives in TMEM columns [288, 320)

struct __align__(128) TcSmem {
    // staging rows, two buffers (sort tile parity) x four TMEM lane quadrants x 32 bodies; the row stride 194 floats = 776 B makes
    // even rows 16 B aligned and odd rows 8 mod 16, exactly like the cloud rows of even / odd bodies (6890 * 12 B = 8 mod 16), so a
    // row leaves as one TMA bulk store (odd bodies: 8 B head and tail by plain stores).  The prologue lands the 128 x 288 joint
    // transforms here first (rows of TC_GROW floats).
    float out[2][4][32 * TC_SROW];
    // B operand ring.  Reference semantics: two stages of a whole K = 32 shape operand (12 KB).  With pose blendshapes: two stages
    // of a K = 16 chunk, [A 128 x 16 | B 96 x 16] = 8 KB + 6 KB (the two shape chunks use the B part only, their A is in TMEM)
    union {
        unsigned char oper[TC_NOP][TC_OPER_BYTES];
        unsigned char chunk[2][TC_CH_BYTES];
    };
    unsigned long long ch_full[2], ch_free[2];
    unsigned char recs[TC_NREC][TC_REC_BYTES];
    unsigned long long op_full[TC_NOP], op_free[TC_NOP], rec_full[TC_NREC], rec_free[TC_NREC], d_full[2], d_empty[2], g_full;
    unsigned long long out_full[2][4], out_free[2][4];   // staging buffer x quadrant: skinning warps -> store warp and back
    uint32_t tmem_base;
};
static_assert(TC_M * TC_GROW * 4 <= (int)sizeof(float) * 2 * 4 * 32 * TC_SROW, "prologue staging of the transforms");
static_assert((sizeof(float) * 2 * 4 * 32 * TC_SROW) % 128 == 0, "operand stage alignment");
static_assert(sizeof(TcSmem) <= 227 * 1024, "shared memory budget");

// canonical K-major, no swizzle (UMMA "interleave"): slabs of 8 k; in a slab [row group of 8][k chunk of 4][row][k]; verified
// against tcgen05.mma on B200 by tools/tc_probe.cu (LBO = 128 B between k chunks, SBO = 256 B between row groups)
__host__ __device__ inline int canon_off(int rows, int r, int k) {
    return (k >> 3) * rows * 8 + (r >> 3) * 64 + ((k >> 2) & 1) * 32 + (r & 7) * 4 + (k & 3);
}

__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s_u32(bar)), "r"(count));
}
__device__ __forceinline__ void bar_expect_tx(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bar_arrive(unsigned long long* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s_u32(bar)) : "memory");
}
__device__ __forceinline__ void bar_wait(unsigned long long* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"   // %2: suspend-time hint, the thread sleeps in hardware
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(s_u32(bar)), "r"(parity), "r"(0x989680u) : "memory");
}
__device__ __forceinline__ bool bar_test(unsigned long long* bar, uint32_t parity) {   // one non-blocking look
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(s_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s_u32(dst)), "l"(src),
                 "r"(bytes), "r"(s_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(unsigned long long* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s_u32(bar)) : "memory");
}
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((128u >> 4) & 0x3FFF) << 16;   // leading-dimension byte offset: between the two k chunks of a K = 8 step
    d |= (uint64_t)((256u >> 4) & 0x3FFF) << 32;   // stride byte offset: between groups of 8 rows
    d |= (uint64_t)1 << 46;                        // descriptor version (sm_100); layout type 0 = no swizzle
    return d;
}
__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
        : "memory");
}
// A operand from TMEM (row = lane, one column per K element), B through a shared-memory descriptor
__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem), "r"(a_tmem), "l"(db), "r"(idesc), "r"(accumulate)
        : "memory");
}
// TMA bulk store shared -> global (both 16 B aligned, bytes % 16 == 0), tracked by the thread's bulk async-group
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(s_u32(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]), "=f"(v[8]), "=f"(v[9]),
                   "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_st4(uint32_t taddr, float4 v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
// packed FP32: one FFMA2 issue slot does two lanes' worth of fused multiply-adds (tools/ffma_probe.cu: the FP32 rate at half the slots)
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ void named_bar(int id, int threads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory"); }

// Optional timeline (debug hook b3d_debug_tc_trace): clock64 of CTA (0, 0)'s warps at five points of every MMA tile.
constexpr int TC_TRACE_TILES = 256, TC_TRACE_PTS = 8;
__device__ long long g_tc_trace[(TC_EPI_WARPS + 4) * TC_TRACE_TILES * TC_TRACE_PTS];

template <bool TRACE, bool POSE>
__global__ void __launch_bounds__(TC_THREADS, 1) smpl_lbs_tc_kernel(const unsigned char* __restrict__ blob, const float* __restrict__ pose_a, const float* __restrict__ pose_b, int V, int B, int sort_tiles, int splits, int bulk_rows, int trace,
                                                                     const float* __restrict__ wshape, const float* __restrict__ G,
                                                                     float* __restrict__ cloud) {
    extern __shared__ __align__(128) unsigned char tc_raw[];
    TcSmem& S = *reinterpret_cast<TcSmem*>(tc_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int b0 = blockIdx.y * TC_M;
    const int t_begin = (int)((long long)blockIdx.x * sort_tiles / splits);
    const int t_end = (int)((long long)(blockIdx.x + 1) * sort_tiles / splits);
    const int nM = TC_MPT * (t_end - t_begin);

    if (tid == 0) {
        for (int i = 0; i < TC_NOP; ++i) { bar_init(&S.op_full[i], 1); bar_init(&S.op_free[i], 1); }
        for (int i = 0; i < TC_NREC; ++i) { bar_init(&S.rec_full[i], 1); bar_init(&S.rec_free[i], TC_EPI_WARPS); }
        for (int i = 0; i < 2; ++i) { bar_init(&S.d_full[i], 1); bar_init(&S.d_empty[i], TC_EPI_WARPS); }
        bar_init(&S.g_full, 1);
        for (int i = 0; i < 2; ++i)
            for (int j = 0; j < 4; ++j) { bar_init(&S.out_full[i][j], TC_WPQ); bar_init(&S.out_free[i][j], 1); }
        for (int i = 0; i < 2; ++i) { bar_init(&S.ch_full[i], 1); bar_init(&S.ch_free[i], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32(&S.tmem_base)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = S.tmem_base;
    const int q = warp & 3, h = warp >> 2;   // h == 4: service warpgroup
    const uint32_t tm_row = tm + ((uint32_t)(32 * q) << 16);
    // joint transforms of the CTA's bodies: one TMA bulk copy per body row (1152 B) into padded shared rows
    float* const gst = &S.out[0][0][0];
    const int nb = min(TC_M, B - b0);
    if (tid == 0) bar_expect_tx(&S.g_full, (uint32_t)nb * NJ * 12 * 4);
    if (tid < nb) bulk_g2s(gst + tid * TC_GROW, G + (size_t)(b0 + tid) * (NJ * 12), NJ * 12 * 4, &S.g_full);
    // A operand -> TMEM columns [288, 320), row = body = lane: TF32 hi / hi / lo copies of the shape vector and two ones for the
    // template columns (k 0-9 w_hi, 10-19 w_hi, 20-29 w_lo, 30-31 one).  tcgen05.mma reads it from TMEM (tools/tc_probe.cu).
    if (h == 0) {
        const int body = b0 + 32 * q + lane;
        float a[TC_K];
#pragma unroll
        for (int s = 0; s < NS; ++s) {
            const float w = body < B ? wshape[(size_t)body * NS + s] : 0.f;
            const float hi = to_tf32(w);
            a[s] = hi; a[NS + s] = hi; a[2 * NS + s] = to_tf32(w - hi);
        }
        a[30] = 1.f; a[31] = 1.f;
#pragma unroll
        for (int c = 0; c < TC_K; c += 4) tmem_st4(tm_row + TC_COL_A + c, make_float4(a[c], a[c + 1], a[c + 2], a[c + 3]));
    }
    // shared rows -> TMEM columns [0, 288), body = lane (rows beyond the batch hold whatever the buffer held: never stored)
    if (h < TC_WPQ) {
        constexpr int CPW = NJ * 12 / TC_WPQ;   // columns per warp
        bar_wait(&S.g_full, 0);
        // per joint the column-major [R | t] is re-ordered to [r00 r10 | r01 r11 | r02 r12 | t0 t1 | r20 r21 r22 t2]: the x / y rows
        // come back from tcgen05.ld as aligned register pairs for the packed FMAs of the slot loop
        const float4* src = reinterpret_cast<const float4*>(gst + (32 * q + lane) * TC_GROW + CPW * h);
#pragma unroll
        for (int j = 0; j < CPW / 12; ++j) {
            const float4 a = src[3 * j], b = src[3 * j + 1], c = src[3 * j + 2];
            tmem_st4(tm_row + CPW * h + 12 * j, make_float4(a.x, a.y, a.w, b.x));
            tmem_st4(tm_row + CPW * h + 12 * j + 4, make_float4(b.z, b.w, c.y, c.z));
            tmem_st4(tm_row + CPW * h + 12 * j + 8, make_float4(a.z, b.y, c.x, c.w));
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // staging rows were read through the generic proxy
    tc_fence_before();
    __syncthreads();   // also: the staging rows may now be overwritten
    tc_fence_after();

    const bool tr = TRACE && trace && blockIdx.x == 0 && blockIdx.y == 0 && lane == 0;
    auto stamp = [&](int m, int k) { if (TRACE && tr && m < TC_TRACE_TILES) g_tc_trace[(warp * TC_TRACE_TILES + m) * TC_TRACE_PTS + k] = clock64(); };
    const unsigned char* const blob0 = blob + (size_t)(TC_MPT * t_begin) * TC_BLOB_BYTES;
    // instruction descriptor: D = F32, A = B = TF32, both K-major, N = 96, M = 128
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TC_N >> 3) << 17) | ((uint32_t)(TC_M >> 4) << 24);
    auto issue_tma = [&](int m) {   // one thread: operand(s) and slot records of MMA tile m -> their rings
        const int sr = m % TC_NREC, ur = m / TC_NREC;
        bar_wait(&S.rec_free[sr], (uint32_t)((ur & 1) ^ 1));       // every warp is done with this record block's last tile
        bar_expect_tx(&S.rec_full[sr], TC_REC_BYTES);
        bulk_g2s(S.recs[sr], blob0 + (size_t)m * TC_BLOB_BYTES + TC_OPER_BYTES, TC_REC_BYTES, &S.rec_full[sr]);
        if (!POSE) {
            const int so = m % TC_NOP, uo = m / TC_NOP;
            bar_wait(&S.op_free[so], (uint32_t)((uo & 1) ^ 1));    // the MMA that read this operand stage is complete
            bar_expect_tx(&S.op_full[so], TC_OPER_BYTES);
            bulk_g2s(S.oper[so], blob0 + (size_t)m * TC_BLOB_BYTES, TC_OPER_BYTES, &S.op_full[so]);
        } else {
            // 15 chunks of K = 16: two of the shape operand (B only), thirteen of the pose contraction (A: this body group's pose
            // features, B: the tile's posedirs columns), through a two-stage ring
            const float* pa = pose_a + (size_t)blockIdx.y * TC_POSE_KSTEPS * (TC_M * 8);
            const float* pb = pose_b + (size_t)(TC_MPT * t_begin + m) * TC_POSE_KSTEPS * (TC_N * 8);
            for (int c = 0; c < TC_NCH; ++c) {
                const int g = m * TC_NCH + c, sc = g & 1, uc = g >> 1;
                bar_wait(&S.ch_free[sc], (uint32_t)((uc & 1) ^ 1));
                if (c < 2) {
                    bar_expect_tx(&S.ch_full[sc], TC_CH_B);
                    bulk_g2s(S.chunk[sc] + TC_CH_A, blob0 + (size_t)m * TC_BLOB_BYTES + (size_t)c * TC_CH_B, TC_CH_B, &S.ch_full[sc]);
                } else {
                    bar_expect_tx(&S.ch_full[sc], TC_CH_BYTES);
                    bulk_g2s(S.chunk[sc], pa + (size_t)(2 * (c - 2)) * (TC_M * 8), TC_CH_A, &S.ch_full[sc]);
                    bulk_g2s(S.chunk[sc] + TC_CH_A, pb + (size_t)(2 * (c - 2)) * (TC_N * 8), TC_CH_B, &S.ch_full[sc]);
                }
            }
        }
    };
    auto issue_mma = [&](int m) {   // one thread: blend(s) of MMA tile m into accumulator buffer m & 1
        const int buf = m & 1, r = m >> 1;
        const uint32_t d_tmem = tm + (uint32_t)(buf ? TC_COL_D1 : TC_COL_D0);
        stamp(m, 0);
        if (!POSE) {
            const int so = m % TC_NOP, uo = m / TC_NOP;
            bar_wait(&S.op_full[so], (uint32_t)(uo & 1));
            bar_wait(&S.d_empty[buf], (uint32_t)((r & 1) ^ 1));
            tc_fence_after();
            stamp(m, 5);
            const uint32_t b_addr = s_u32(S.oper[so]);
#pragma unroll
            for (int s = 0; s < TC_K / 8; ++s)
                mma_tf32_ts(d_tmem, tm + (uint32_t)(TC_COL_A + 8 * s), smem_desc(b_addr + s * TC_B_SLAB), idesc, s > 0 ? 1u : 0u);
            stamp(m, 6);
            tc_commit(&S.op_free[so]);
        } else {
            bar_wait(&S.d_empty[buf], (uint32_t)((r & 1) ^ 1));
            for (int c = 0; c < TC_NCH; ++c) {
                const int g = m * TC_NCH + c, sc = g & 1, uc = g >> 1;
                bar_wait(&S.ch_full[sc], (uint32_t)(uc & 1));
                tc_fence_after();
                const uint32_t a_addr = s_u32(S.chunk[sc]), b_addr = a_addr + TC_CH_A;
#pragma unroll
                for (int s = 0; s < 2; ++s) {
                    if (c < 2) mma_tf32_ts(d_tmem, tm + (uint32_t)(TC_COL_A + 16 * c + 8 * s), smem_desc(b_addr + s * TC_B_SLAB), idesc, (c | s) ? 1u : 0u);
                    else mma_tf32(d_tmem, smem_desc(a_addr + s * TC_A_SLAB), smem_desc(b_addr + s * TC_B_SLAB), idesc, 1u);
                }
                tc_commit(&S.ch_free[sc]);
            }
        }
        tc_commit(&S.d_full[buf]);
        stamp(m, 7);
    };
    if (warp >= TC_EPI_WARPS) {
        // service warpgroup: hands most of its registers to the skinning warps, then two of its threads run the copy / MMA loops
        asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
        if (warp == TC_EPI_WARPS && lane == 0)
            for (int m = 0; m < nM; ++m) issue_tma(m);
        if (warp == TC_EPI_WARPS + 1 && lane == 0)
            for (int m = 0; m < nM; ++m) issue_mma(m);
        if (warp >= TC_EPI_WARPS + 2 && bulk_rows) {
            // store warps (18: quadrants 0 and 2, 19: quadrants 1 and 3), lane = body of the quadrant.  Issuing a bulk store blocks
            // the issuing warp until the SM's store path takes it (a quadrant's 24 KB: ~1,300 cycles, all four skinning warps of a
            // scheduler at once when they issued their own rows) - here it blocks a warp that has nothing else to do.
            const int nfull = min(t_end, V / TC_T) - t_begin;   // complete sort tiles (the mesh's last, partial one leaves by plain stores)
            const int q0 = warp - (TC_EPI_WARPS + 2);
            int sent[2] = {0, 0};
            while (sent[0] < nfull || sent[1] < nfull) {
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int st = sent[i], qq = q0 + 2 * i, buf = st & 1;
                    if (st >= nfull) continue;
                    if (!__all_sync(0xffffffffu, bar_test(&S.out_full[buf][qq], (uint32_t)((st >> 1) & 1)))) continue;
                    const int body = b0 + 32 * qq + lane;
                    if (body < B) {
                        const float* sp = &S.out[buf][qq][lane * TC_SROW];
                        float* dp = cloud + ((size_t)body * V + (size_t)(t_begin + st) * TC_T) * 3;
                        if ((reinterpret_cast<uintptr_t>(dp) & 15) == 0) {
                            bulk_s2g(dp, sp, 3 * TC_T * 4);
                        } else {   // 8 mod 16, as is the shared row: head and tail pairs by hand, the 16 B aligned middle in bulk
                            *reinterpret_cast<float2*>(dp) = *reinterpret_cast<const float2*>(sp);
                            *reinterpret_cast<float2*>(dp + 3 * TC_T - 2) = *reinterpret_cast<const float2*>(sp + 3 * TC_T - 2);
                            bulk_s2g(dp + 2, sp + 2, 3 * TC_T * 4 - 16);
                        }
                    }
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the rows have left shared memory
                    __syncwarp();
                    if (lane == 0) bar_arrive(&S.out_free[buf][qq]);
                    sent[i] = st + 1;
                }
            }
            asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");
        float g[4][12];
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int e = 0; e < 12; ++e) g[k][e] = 0.f;
        constexpr int RPW = 32 / TC_WPQ;   // staging rows a warp sends off (plain-store path)
        for (int m = 0; m < nM; ++m) {
            stamp(m, 0);
            const int sr = m % TC_NREC, ur = m / TC_NREC, buf = m & 1, r = m >> 1;
            const int st_tile = m / TC_MPT;                    // sort tile (local); its staging buffer alternates
            const uint32_t orow_s = s_u32(&S.out[st_tile & 1][q][lane * TC_SROW]);
            // the store warp has sent off the rows this staging buffer held two sort tiles ago
            if (bulk_rows && m % TC_MPT == 0 && st_tile >= 2) bar_wait(&S.out_free[st_tile & 1][q], (uint32_t)(((st_tile >> 1) - 1) & 1));
            bar_wait(&S.rec_full[sr], (uint32_t)(ur & 1));     // slot records of this tile are in shared memory
            bar_wait(&S.d_full[buf], (uint32_t)(r & 1));       // and its accumulators are complete
            tc_fence_after();
            stamp(m, 1);
            float xyz[3 * TC_SPW];   // accumulators of this warp's slots
            const uint32_t d_addr = tm_row + (uint32_t)(buf ? TC_COL_D1 : TC_COL_D0) + (uint32_t)(3 * TC_SPW * h);
#pragma unroll
            for (int c = 0; c < 3 * TC_SPW; c += 8) tmem_ld8(d_addr + c, xyz + c);
            tc_wait_ld();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) bar_arrive(&S.d_empty[buf]);        // the accumulator buffer may be overwritten
            stamp(m, 2);

            const uint4* rec = reinterpret_cast<const uint4*>(S.recs[sr]) + 2 * (TC_SPW * h);
            uint4 rq[TC_SPW + 1][2];   // statically indexed (the loop is unrolled): the next slot's record hides behind this slot
            rq[0][0] = rec[0]; rq[0][1] = rec[1];
#pragma unroll
            for (int s = 0; s < TC_SPW; ++s) {
                if (s + 1 < TC_SPW) { rq[s + 1][0] = rec[2 * s + 2]; rq[s + 1][1] = rec[2 * s + 3]; }
                const uint4 r0 = rq[s][0], r1 = rq[s][1];
                uint32_t ctl = r1.y;
                if (s == 0 && m == 0) ctl |= 15u;   // the CTA's first slot: nothing is in registers yet (the control word names all four joints)
                if (ctl & 15u) {   // warp-uniform: some of the four register sets change joint
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (ctl & (1u << k)) {
                            const uint32_t col = 12u * ((ctl >> (4 + 5 * k)) & 31u);
                            tmem_ld8(tm_row + col, g[k]);
                            tmem_ld4(tm_row + col + 8, g[k] + 8);
                        }
                    tc_wait_ld();
                }
                const float x = xyz[3 * s], y = xyz[3 * s + 1], z = xyz[3 * s + 2];
                // x and y rows: one packed FMA chain per joint, sum_k w_k (R_k p + t_k) in that order; z row: the weighted row
                // sum_k w_k [r20 r21 r22 t2]_k first (two packed FMAs per joint), then one dot product.  27 issue slots instead of 48.
                const unsigned long long X = pack2(x, x), Y = pack2(y, y), Z = pack2(z, z);
                const float wk[4] = {__uint_as_float(r0.x), __uint_as_float(r0.y), __uint_as_float(r0.z), __uint_as_float(r0.w)};
                unsigned long long axy = 0ull, mz01 = 0ull, mz23 = 0ull;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const unsigned long long W = pack2(wk[k], wk[k]);   // ptxas folds (v, v) pairs into the FFMA2 broadcast operand form
                    const unsigned long long txy = fma2(pack2(g[k][0], g[k][1]), X, fma2(pack2(g[k][2], g[k][3]), Y, fma2(pack2(g[k][4], g[k][5]), Z, pack2(g[k][6], g[k][7]))));
                    axy = fma2(W, txy, axy);
                    mz01 = fma2(W, pack2(g[k][8], g[k][9]), mz01);
                    mz23 = fma2(W, pack2(g[k][10], g[k][11]), mz23);
                }
                float ax, ay, m0, m1, m2, m3;
                unpack2(axy, ax, ay);
                unpack2(mz01, m0, m1);
                unpack2(mz23, m2, m3);
                const float az = fmaf(m0, x, fmaf(m1, y, fmaf(m2, z, m3)));
                // no "memory" clobber: the stores only have to stay ahead of the (volatile, clobbering) barrier arrive below
                const uint32_t so = orow_s + (r1.x << 2);
                asm volatile("st.shared.f32 [%0], %1;\n\tst.shared.f32 [%0+4], %2;\n\tst.shared.f32 [%0+8], %3;" ::"r"(so), "f"(ax), "f"(ay), "f"(az));
            }
            __syncwarp();
            if (lane == 0) bar_arrive(&S.rec_free[sr]);        // records read: the block may be refilled
            stamp(m, 3);

            if (m % TC_MPT == TC_MPT - 1) {   // a sort tile is complete: rows of 3 x 64 consecutive floats per body
                const int base_v = (t_begin + st_tile) * TC_T;
                const int nv = min(TC_T, V - base_v);
                if (nv == TC_T && bulk_rows) {   // hand the quadrant's rows to the store warp
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // this tile's staging stores -> visible to the bulk copies
                    __syncwarp();
                    if (lane == 0) bar_arrive(&S.out_full[st_tile & 1][q]);
                } else {   // the mesh's last, partial sort tile (or a cloud the bulk stores cannot address): plain stores
                    named_bar(1 + q, 32 * TC_WPQ);
                    const int row0 = RPW * h;                                   // this warp's rows of the quadrant
                    const int rows = min(RPW, B - (b0 + 32 * q + row0));
                    const float* srow = &S.out[st_tile & 1][q][row0 * TC_SROW];
                    float* drow = cloud + ((size_t)(b0 + 32 * q + row0) * V + base_v) * 3;
                    const int nfl = 3 * nv;
                    for (int row = 0; row < rows; ++row)
                        for (int e = lane; e < nfl; e += 32) drow[(size_t)row * V * 3 + e] = srow[row * TC_SROW + e];
                }
            }
            stamp(m, 4);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(512));
}

inline float tf32_round_host(float x) {   // round to 10 explicit mantissa bits (ties away); the low 13 bits end up zero
    uint32_t u;
    std::memcpy(&u, &x, 4);
    u = (u + 0x1000u) & ~0x1FFFu;
    float r;
    std::memcpy(&r, &u, 4);
    return r;
}

}  // namespace

// Host side of the tables: slot order, register-set assignment and the operand blobs.
int build_tc_tables(b3d_model& m, const std::vector<double>& vert, const std::vector<double>& wval, const std::vector<int32_t>& wjoint) {
    const size_t Vp = (size_t)m.Vpad;
    const int V = m.V;
    const int tiles = (V + TC_T - 1) / TC_T;
    std::vector<unsigned char> blob((size_t)tiles * TC_MPT * TC_BLOB_BYTES, 0);
    m.tc.slot_vertex.assign((size_t)tiles * TC_MPT * TC_SLOTS, -1);
    long long reloads = 0, slots_total = 0;

    struct Vtx { int v; int nj; int j[4]; float w[4]; uint64_t key; };
    // register sets of each slot sequence (= warp of a quadrant), carried from sort tile to sort tile: a warp walks the tiles in
    // order, and a CTA that starts in the middle re-reads all four sets at its first slot
    int reg_state[TC_SEQS][4];
    for (int h = 0; h < TC_SEQS; ++h)
        for (int k = 0; k < 4; ++k) reg_state[h][k] = -1;
    for (int t = 0; t < tiles; ++t) {
        const int base = t * TC_T;
        std::vector<Vtx> vs(TC_T);
        int pad_pos = std::min(TC_T, V - base);
        for (int i = 0; i < TC_T; ++i) {
            Vtx& x = vs[i];
            x.v = base + i;
            x.nj = 0;
            x.key = ~0ull;
            if (x.v < V) {
                uint64_t k = 0;
                for (int q = 0; q < 4; ++q) {
                    const double wv = wval[q * Vp + x.v];
                    if (wv != 0.0) { x.j[x.nj] = wjoint[q * Vp + x.v]; x.w[x.nj] = (float)wv; ++x.nj; }
                }
                for (int q = 0; q < 4; ++q) k = k * 32 + (uint64_t)(q < x.nj ? x.j[q] : 31);
                x.key = k;
            }
        }
        std::vector<int> order(TC_T);
        for (int i = 0; i < TC_T; ++i) order[i] = i;
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return vs[a].key < vs[b].key; });

        constexpr int SEQ = TC_T / TC_SEQS, SPW = TC_SLOTS / TC_SEQS;   // slots per sequence in a sort tile / in an MMA tile
        for (int h = 0; h < TC_SEQS; ++h) {
            // the sequence's vertices, grouped by joint set, groups ordered greedily by the number of register sets to re-read
            std::vector<int> idx(order.begin() + SEQ * h, order.begin() + SEQ * h + SEQ);
            std::vector<std::vector<int>> groups;
            for (int i : idx) {
                if (!groups.empty() && vs[groups.back()[0]].key == vs[i].key) groups.back().push_back(i);
                else groups.push_back({i});
            }
            int* reg = reg_state[h];
            std::vector<bool> used(groups.size(), false);
            int slot = 0;
            for (size_t done = 0; done < groups.size(); ++done) {
                int best = -1, best_cost = 99;
                for (size_t gi = 0; gi < groups.size(); ++gi) {
                    if (used[gi]) continue;
                    const Vtx& x = vs[groups[gi][0]];
                    int cost = 0;
                    for (int a = 0; a < x.nj; ++a) cost += !(reg[0] == x.j[a] || reg[1] == x.j[a] || reg[2] == x.j[a] || reg[3] == x.j[a]);
                    if (done == 0 && reg[0] < 0) cost = (int)gi;   // nothing loaded yet: start from the first group in key order
                    if (cost < best_cost) { best_cost = cost; best = (int)gi; }
                }
                used[best] = true;
                const Vtx& g0 = vs[groups[best][0]];
                // assign missing joints to register sets that hold joints this group does not use
                uint32_t flags = 0;
                for (int a = 0; a < g0.nj; ++a) {
                    bool have = false;
                    for (int k = 0; k < 4; ++k) have = have || reg[k] == g0.j[a];
                    if (have) continue;
                    // a set is taken if it is the FIRST one that holds a joint of this group (the weights below go to the first
                    // match); a second copy of such a joint - the unused sets start out on joint 0 - is free like any other
                    int k_free = -1;
                    for (int k = 0; k < 4 && k_free < 0; ++k) {
                        bool needed = false;
                        for (int a2 = 0; a2 < g0.nj; ++a2) needed = needed || reg[k] == g0.j[a2];
                        for (int k2 = 0; k2 < k; ++k2) needed = needed && reg[k2] != reg[k];
                        if (!needed) k_free = k;
                    }
                    if (k_free < 0) return B3D_EINVAL;   // cannot happen: at most three of the four sets are first holders of other joints
                    reg[k_free] = g0.j[a];
                    flags |= 1u << k_free;
                }
                for (size_t gi2 = 0; gi2 < groups[best].size(); ++gi2, ++slot) {
                    const Vtx& x = vs[groups[best][gi2]];
                    uint32_t f = gi2 == 0 ? flags : 0u;
                    int jr[4];
                    for (int k = 0; k < 4; ++k) {
                        if (reg[k] < 0) { reg[k] = 0; f |= 1u << k; }   // very first slot: unused sets hold joint 0
                        jr[k] = reg[k];
                    }
                    const int mt = slot / SPW, s = slot % SPW;
                    unsigned char* bl = &blob[((size_t)t * TC_MPT + mt) * TC_BLOB_BYTES];
                    float* oper = reinterpret_cast<float*>(bl);
                    uint32_t* rec = reinterpret_cast<uint32_t*>(bl + TC_OPER_BYTES) + (SPW * h + s) * TC_REC_WORDS;
                    float wr[4] = {0.f, 0.f, 0.f, 0.f};
                    for (int a = 0; a < x.nj; ++a)
                        for (int k = 0; k < 4; ++k)
                            if (jr[k] == x.j[a]) { wr[k] = x.w[a]; break; }
                    std::memcpy(rec, wr, 16);
                    int pos = x.v < V ? x.v - base : pad_pos++;
                    rec[4] = (uint32_t)(3 * pos);
                    rec[5] = f | ((uint32_t)jr[0] << 4) | ((uint32_t)jr[1] << 9) | ((uint32_t)jr[2] << 14) | ((uint32_t)jr[3] << 19);
                    rec[6] = (uint32_t)x.v;
                    rec[7] = 0;
                    if (x.v < V) m.tc.slot_vertex[((size_t)t * TC_MPT + mt) * TC_SLOTS + SPW * h + s] = x.v;
                    for (int k = 0; k < 4; ++k) reloads += (f >> k) & 1u;
                    ++slots_total;
                    if (x.v < V)
                        for (int c = 0; c < 3; ++c) {
                            const int n = 3 * (SPW * h + s) + c;
                            for (int k = 0; k < NS; ++k) {
                                const float sd = (float)vert[(3 + NS * c + k) * Vp + x.v];
                                const float hi = tf32_round_host(sd), lo = tf32_round_host(sd - hi);
                                oper[canon_off(TC_N, n, k)] = hi;
                                oper[canon_off(TC_N, n, NS + k)] = lo;
                                oper[canon_off(TC_N, n, 2 * NS + k)] = hi;
                            }
                            const float tv = (float)vert[c * Vp + x.v];
                            const float hi = tf32_round_host(tv), lo = tf32_round_host(tv - hi);
                            oper[canon_off(TC_N, n, 30)] = hi;
                            oper[canon_off(TC_N, n, 31)] = lo;
                        }
                }
            }
        }
    }
    B3D_CUDA(cudaMalloc(&m.tc.blob, blob.size()));
    B3D_CUDA(cudaMemcpy(m.tc.blob, blob.data(), blob.size(), cudaMemcpyHostToDevice));
    m.tc.sort_tiles = tiles;
    m.tc.mean_reloads = slots_total ? (double)reloads / (double)slots_total : 0.0;
    return B3D_OK;
}

static std::atomic<int> g_tc_trace_on{0};
static std::atomic<int> g_tc_forced_splits{0};

// Pose features (B x 207 row-major, joints kernel) -> A operand slabs of the pose contraction: per group of 128 bodies, 26 slabs
// [128 rows x 8 k] in the canonical layout, rounded to TF32 (one pass is enough: the pose term is ~1e-2 of a vertex coordinate).
__global__ void __launch_bounds__(256) pose_operand_kernel(const float* __restrict__ pf, int B, float* __restrict__ pose_a) {
    const int g = blockIdx.x;
    for (int i = threadIdx.x; i < TC_M * 8 * TC_POSE_KSTEPS; i += 256) {
        const int r = i / (8 * TC_POSE_KSTEPS), k = i - r * (8 * TC_POSE_KSTEPS);   // consecutive threads: consecutive k of one body
        const int body = g * TC_M + r;
        const float v = (body < B && k < B3D_POSEDIRS_DIM) ? pf[(size_t)body * B3D_POSEDIRS_DIM + k] : 0.f;
        pose_a[((size_t)g * TC_POSE_KSTEPS + (k >> 3)) * (TC_M * 8) + canon_off(TC_M, r, k & 7)] = to_tf32(v);
    }
}

int launch_lbs_tc(const b3d_model* m, int B, const float* wshape, const float* G, const float* posefeat, float* cloud, cudaStream_t st) {
    // the attribute is per device and context: set it for the instantiation this call launches (cheap), not once per process
    const int trace = g_tc_trace_on.load();
    {
        const void* fn = posefeat ? (const void*)smpl_lbs_tc_kernel<false, true> : trace ? (const void*)smpl_lbs_tc_kernel<true, false>
                                                                                        : (const void*)smpl_lbs_tc_kernel<false, false>;
        B3D_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TcSmem)));
    }
    const int groups = (B + TC_M - 1) / TC_M;
    const int tiles = m->tc.sort_tiles;
    // vertex splits: fill 148 SMs in whole waves; a CTA's prologue (TMEM allocation, transforms -> TMEM, pipeline fill) costs about
    // one sort tile's worth of time (fitted to B = 65536 timings at 2 and 8 splits)
    const int forced = g_tc_forced_splits.load();   // debug hook b3d_debug_tc_splits (tests vary the vertex split); 0 in production
    int best_s = 1;
    double best_cost = 1e30;
    for (int s = 1; s <= std::min(tiles, 24); ++s) {
        const long long ctas = (long long)groups * s;
        const long long waves = (ctas + 147) / 148;
        const int per = (tiles + s - 1) / s;   // sort tiles of the longest CTA
        const double cost = (double)waves * ((double)per + 1.1);
        if (cost < best_cost * 0.999) { best_cost = cost; best_s = s; }
    }
    if (forced > 0) best_s = std::min(forced, tiles);
    // rows leave as TMA bulk stores when row r of the cloud is 16 B aligned for even r and 8 mod 16 for odd r (true for V = 6890 and
    // a 16 B aligned cloud: 6890 * 12 = 8 mod 16), which is how the staging rows are laid out; otherwise plain stores
    const int bulk_rows = ((reinterpret_cast<uintptr_t>(cloud) & 15) == 0 && ((size_t)m->V * 12) % 16 == 8) ? 1 : 0;
    const dim3 grid(best_s, groups);
    if (posefeat) {
        if (!m->tc.pose_b) return set_error(B3D_EINVAL, "tensor-core LBS: the model has no pose blendshape tables");
        float* pose_a = nullptr;
        B3D_CUDA(cudaMallocAsync(&pose_a, sizeof(float) * (size_t)groups * TC_POSE_KSTEPS * TC_M * 8, st));
        pose_operand_kernel<<<groups, 256, 0, st>>>(posefeat, B, pose_a);
        B3D_LAUNCH_CHECK();
        smpl_lbs_tc_kernel<false, true><<<grid, TC_THREADS, sizeof(TcSmem), st>>>(m->tc.blob, pose_a, m->tc.pose_b, m->V, B, tiles, best_s, bulk_rows, 0, wshape, G, cloud);
        B3D_CUDA(cudaFreeAsync(pose_a, st));
    } else if (trace) {
        smpl_lbs_tc_kernel<true, false><<<grid, TC_THREADS, sizeof(TcSmem), st>>>(m->tc.blob, nullptr, nullptr, m->V, B, tiles, best_s, bulk_rows, trace, wshape, G, cloud);
    } else {
        smpl_lbs_tc_kernel<false, false><<<grid, TC_THREADS, sizeof(TcSmem), st>>>(m->tc.blob, nullptr, nullptr, m->V, B, tiles, best_s, bulk_rows, 0, wshape, G, cloud);
    }
    return B3D_OK;
}

// B operand slabs of the pose contraction, in every MMA tile's slot order (p32[(k * 3 + c) * Vpad + v], model.cu)
int build_tc_pose_tables(b3d_model& m, const std::vector<float>& p32) {
    if (m.tc.slot_vertex.empty()) return B3D_OK;
    const size_t Vp = (size_t)m.Vpad;
    const int mtiles = m.tc.sort_tiles * TC_MPT;
    std::vector<float> pb((size_t)mtiles * TC_POSE_KSTEPS * TC_N * 8, 0.f);
    for (int mt = 0; mt < mtiles; ++mt)
        for (int slot = 0; slot < TC_SLOTS; ++slot) {
            const int v = m.tc.slot_vertex[(size_t)mt * TC_SLOTS + slot];
            if (v < 0) continue;
            for (int c = 0; c < 3; ++c)
                for (int k = 0; k < B3D_POSEDIRS_DIM; ++k)
                    pb[((size_t)mt * TC_POSE_KSTEPS + (k >> 3)) * (TC_N * 8) + canon_off(TC_N, 3 * slot + c, k & 7)] =
                        tf32_round_host(p32[((size_t)k * 3 + c) * Vp + v]);
        }
    B3D_CUDA(cudaMalloc(&m.tc.pose_b, sizeof(float) * pb.size()));
    B3D_CUDA(cudaMemcpy(m.tc.pose_b, pb.data(), sizeof(float) * pb.size(), cudaMemcpyHostToDevice));
    return B3D_OK;
}

}  // namespace b3d

extern "C" {
// debug hooks (tools/prof.py tctrace): switch the timeline on / read it back as [warp][tile][point] clock64 values
int b3d_debug_tc_trace(int on) { b3d::g_tc_trace_on.store(on); return 0; }
// force the number of vertex splits of the tensor-core LBS launch (0 = automatic); tests use it to cover every tile split
int b3d_debug_tc_splits(int splits) { b3d::g_tc_forced_splits.store(splits > 0 ? splits : 0); return 0; }
int b3d_debug_tc_trace_read(long long* out, int n) {
    const size_t bytes = sizeof(long long) * (size_t)std::min(n, (b3d::TC_EPI_WARPS + 4) * b3d::TC_TRACE_TILES * b3d::TC_TRACE_PTS);
    return cudaMemcpyFromSymbol(out, b3d::g_tc_trace, bytes) == cudaSuccess ? 0 : 2;
}
}
