// Flash attention on tcgen05 / TMEM for sm_100a.
//
// One CTA owns 128 query rows of one (batch, head) and streams K/V in tiles of 128 keys:
//   warp 0      TMA producer: Q once, then K_j / V_j into a ring of shared-memory stages
//   warp 1      owns TMEM, issues S_j = Q K_j^T (SS MMA, fp32 in TMEM, double buffered) and
//               O += P_j V_j (P from shared memory K-major, V consumed MN-major as TMA wrote it)
//   warps 2..5  softmax: one thread per query row, tcgen05.ld of the S row, online max/sum in
//               fp32, P written to shared memory in the 128-byte-swizzled UMMA layout
// O lives in TMEM for the whole KV loop. The running max is only raised when it grows by more
// than 2^8 (lazy rescale), so the O read-modify-write through tcgen05.ld/st is rare.
// Token-major strided layout: q/k/v/out are [B, L, H, D] views of the projection outputs, so no
// permute/contiguous copies exist anywhere around the kernel. Tile order over K/V is fixed, so
// the result is bit-reproducible; the multi-GPU variant (parallel/patch_attention) walks the same
// tiles in the same order with some of them living in peer memory.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "../include/arbius_b200.h"
#include "common.cuh"

extern "C" void ab_count_launch(int n);
extern "C" int ab_make_tensor_map(CUtensorMap* out, int dtype, int rank, const void* base, const uint64_t* dims,
                                  const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes);

namespace ab {

constexpr int ATT_BM = 128;   // query rows per CTA
constexpr int ATT_BN = 128;   // keys per tile
constexpr int ATT_THREADS = 192;
constexpr int CHUNK_BYTES = 128 * 128;   // 128 rows x 64 elements x 2 B

struct AttnParams {
    void* out;
    int batch, heads, lq, lk, head_dim;
    long long o_bs, o_ls;
    float scale_log2;   // softmax scale * log2(e)
    int causal;
    int tiles_per_src;  // multi-source K/V: 128-key tiles held by each source (patch-parallel ranks)
    // segmented keys [prefix | pixels] (SegMaps): the prefix fills its own tiles, pixel tiles count from pixel 0 and may
    // straddle sources; K / V of head h are heads pk0 + h / pv0 + h of the prefix buffer, k0 + h / v0 + h of a pixel buffer
    int prefix_len, prefix_tiles, pix_len, rows_per_src, nsrc;
    int pk0, pv0, k0, v0;
    const unsigned int* tick;        // cross-GPU flags: wait for source s (s != self_rank) before the first load from it
    const unsigned int* wait_site;
    int self_rank;
    long long* dbg;                  // profiling aid (ab_attn_set_debug): CTA (0,0,0) of the v3 kernel records a timeline
};

static long long* g_attn_dbg = nullptr;

template <int NCH, typename T>
__global__ void __launch_bounds__(ATT_THREADS, 1)
flash_attn_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                  const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
    constexpr int STAGES = NCH >= 3 ? 1 : 2;
    constexpr int PBUF = NCH >= 3 ? 1 : 2;
    constexpr int QB = NCH * CHUNK_BYTES;          // Q tile bytes, also K or V bytes per stage
    constexpr uint32_t S_COL0 = 0, S_COL1 = 128, O_COL = 256;
    constexpr float kRescaleThreshold = 8.0f;      // log2 units

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + QB;
    uint8_t* sV = sK + STAGES * QB;
    uint8_t* sP = sV + STAGES * QB;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sP + PBUF * 2 * CHUNK_BYTES);
    uint64_t* q_full = bars;                 // 1
    uint64_t* kv_full = bars + 1;            // STAGES
    uint64_t* kv_empty = kv_full + STAGES;   // STAGES
    uint64_t* s_full = kv_empty + STAGES;    // 2
    uint64_t* p_full = s_full + 2;           // PBUF
    uint64_t* o_full = p_full + PBUF;        // 1
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * ATT_BM;
    const int h = blockIdx.y;
    const int b = blockIdx.z;
    const int D = p.head_dim;
    const int DP = (D + 15) & ~15;                       // MMA K (QK^T) and N (PV) extent
    int kv_len = p.lk;
    if (p.causal) kv_len = min(p.lk, q0 + ATT_BM);       // keys beyond the last query row are never needed
    const int NT = (kv_len + ATT_BN - 1) / ATT_BN;   // KV tiles

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
        mbar_init(q_full, 1);
        for (int s = 0; s < STAGES; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
        mbar_init(&s_full[0], 1); mbar_init(&s_full[1], 1);
        for (int s = 0; s < PBUF; ++s) mbar_init(&p_full[s], 128);
        mbar_init(o_full, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch_dependents();

    if (warp == 0) {
        if (elect_one()) {   // single-lane region the compiler can see: operands stay in uniform registers
            mbar_expect_tx(q_full, QB);
#pragma unroll
            for (int c = 0; c < NCH; ++c) tma_load_4d(sQ + c * CHUNK_BYTES, &tmQ, q_full, c * 64, h, q0, b);
            int stage = 0; uint32_t phase = 0;
            for (int j = 0; j < NT; ++j) {
                mbar_wait(&kv_empty[stage], phase ^ 1);
                mbar_expect_tx(&kv_full[stage], 2 * QB);
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    tma_load_4d(sK + stage * QB + c * CHUNK_BYTES, &tmK, &kv_full[stage], c * 64, h, j * ATT_BN, b);
                    tma_load_4d(sV + stage * QB + c * CHUNK_BYTES, &tmV, &kv_full[stage], c * 64, h, j * ATT_BN, b);
                }
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {   // single-lane region the compiler can see: operands stay in uniform registers
            const uint32_t idesc_qk = make_idesc_f16(ATT_BM, ATT_BN, DT<T>::kBf16, false, false);
            const uint32_t idesc_pv = make_idesc_f16(ATT_BM, DP, DT<T>::kBf16, false, true);
            const uint32_t q_addr = smem_u32(sQ);
            auto issue_qk = [&](int j) {
                const int st = j % STAGES;
                mbar_wait(&kv_full[st], (j / STAGES) & 1);
                tc_fence_after();
                const uint32_t k_addr = smem_u32(sK + st * QB);
                const uint32_t d_tmem = tmem_base + ((j & 1) ? S_COL1 : S_COL0);
                for (int ks = 0; ks < DP / 16; ++ks) {
                    const uint32_t off = (ks >> 2) * CHUNK_BYTES + (ks & 3) * 32;
                    umma_ss(d_tmem, make_smem_desc(q_addr + off, 16, 1024, kSwizzle128),
                            make_smem_desc(k_addr + off, 16, 1024, kSwizzle128), idesc_qk, ks > 0 ? 1u : 0u);
                }
                umma_commit(&s_full[j & 1]);
            };
            auto issue_pv = [&](int j) {
                const int st = j % STAGES;
                const int pb = j % PBUF;
                mbar_wait(&p_full[pb], (j / PBUF) & 1);
                tc_fence_after();
                const uint32_t p_addr = smem_u32(sP + pb * 2 * CHUNK_BYTES);
                const uint32_t v_addr = smem_u32(sV + st * QB);
                for (int ks = 0; ks < ATT_BN / 16; ++ks) {
                    const uint64_t adesc = make_smem_desc(p_addr + (ks >> 2) * CHUNK_BYTES + (ks & 3) * 32, 16, 1024, kSwizzle128);
                    // V tile as TMA wrote it: row = key, 64 head-dim elements per 128-byte row, one
                    // 16 KB slab per 64-wide head-dim chunk -> MN-major: LBO = slab stride, SBO = 8 rows.
                    const uint64_t bdesc = make_smem_desc(v_addr + ks * 16 * 128, CHUNK_BYTES, 1024, kSwizzle128);
                    umma_ss(tmem_base + O_COL, adesc, bdesc, idesc_pv, (j > 0 || ks > 0) ? 1u : 0u);
                }
                umma_commit(&kv_empty[st]);
                umma_commit(o_full);
            };
            mbar_wait(q_full, 0);
            if (STAGES >= 2) {
                issue_qk(0);
                for (int j = 0; j < NT; ++j) {
                    if (j + 1 < NT) issue_qk(j + 1);
                    issue_pv(j);
                }
            } else {
                for (int j = 0; j < NT; ++j) { issue_qk(j); issue_pv(j); }
            }
        }
    } else {
        // ================= softmax + epilogue: one thread per query row =================
        const int qd = warp & 3;
        const int r = qd * 32 + lane;
        const int qrow = q0 + r;
        const uint32_t lane_off = static_cast<uint32_t>(qd * 32) << 16;
        float m_ref = -INFINITY;   // reference max (log2 domain) used for the stored exponentials
        float l_sum = 0.f;
        for (int j = 0; j < NT; ++j) {
            mbar_wait(&s_full[j & 1], (j >> 1) & 1);
            tc_fence_after();
            const uint32_t s_addr = tmem_base + lane_off + ((j & 1) ? S_COL1 : S_COL0);
            const int col0 = j * ATT_BN;
            int col_hi = p.lk;                                    // first masked column
            if (p.causal) col_hi = min(col_hi, qrow + 1);
            // ---- pass 1: row max ----
            float mx = -INFINITY;
#pragma unroll 1
            for (int c = 0; c < ATT_BN / 32; ++c) {
                uint32_t raw[32];
                tmem_ld32(s_addr + c * 32, raw);
                tmem_wait_ld();
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float s = __uint_as_float(raw[i]) * p.scale_log2;
                    if (col0 + c * 32 + i < col_hi) mx = fmaxf(mx, s);
                }
            }
            bool need = false;
            float factor = 1.f;
            if (mx > m_ref + kRescaleThreshold || m_ref == -INFINITY) {
                if (mx != -INFINITY) {
                    need = (m_ref != -INFINITY);
                    factor = need ? exp2f(m_ref - mx) : 1.f;
                    m_ref = mx;
                }
            }
            const bool warp_need = __any_sync(0xffffffffu, need);
            if (j > 0) {
                mbar_wait(o_full, (j - 1) & 1);     // PV(j-1) done: O stable, P buffer reusable
                tc_fence_after();
            }
            if (warp_need) {
                l_sum *= factor;
                const uint32_t o_addr = tmem_base + lane_off + O_COL;
#pragma unroll 1
                for (int c = 0; c < DP; c += 16) {
                    uint32_t raw[16];
                    tmem_ld16(o_addr + c, raw);
                    tmem_wait_ld();
#pragma unroll
                    for (int i = 0; i < 16; ++i) raw[i] = __float_as_uint(__uint_as_float(raw[i]) * factor);
                    asm volatile(
                        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, "
                        "%12, %13, %14, %15, %16};\n" ::"r"(o_addr + c),
                        "r"(raw[0]), "r"(raw[1]), "r"(raw[2]), "r"(raw[3]), "r"(raw[4]), "r"(raw[5]), "r"(raw[6]),
                        "r"(raw[7]), "r"(raw[8]), "r"(raw[9]), "r"(raw[10]), "r"(raw[11]), "r"(raw[12]), "r"(raw[13]),
                        "r"(raw[14]), "r"(raw[15])
                        : "memory");
                }
                tmem_wait_st();
            }
            // ---- pass 2: exponentials, row sum, P -> shared memory (K-major, 128B swizzle) ----
            uint8_t* pbuf = sP + (j % PBUF) * 2 * CHUNK_BYTES;
#pragma unroll 1
            for (int c = 0; c < ATT_BN / 32; ++c) {
                uint32_t raw[32];
                tmem_ld32(s_addr + c * 32, raw);
                tmem_wait_ld();
                float pv[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float s = __uint_as_float(raw[i]) * p.scale_log2;
                    const float e = (col0 + c * 32 + i < col_hi) ? exp2f(s - m_ref) : 0.f;
                    pv[i] = e;
                }
                // sum what the tensor core will actually see (values rounded to 16 bits)
                uint8_t* slab = pbuf + (c >> 1) * CHUNK_BYTES + r * 128;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    float f8[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) f8[i] = pv[8 * g + i];
                    const uint4 packed = pack8<T>(f8);
                    float back[8];
                    unpack8<T>(packed, back);
#pragma unroll
                    for (int i = 0; i < 8; ++i) l_sum += back[i];
                    const int chunk16 = (c & 1) * 4 + g;              // 16-byte chunk inside the 128-byte row
                    *reinterpret_cast<uint4*>(slab + ((chunk16 ^ (r & 7)) << 4)) = packed;
                }
            }
            fence_proxy_async();
            tc_fence_before();
            mbar_arrive(&p_full[j % PBUF]);
        }
        // ---- epilogue: O / l -> out ----
        mbar_wait(o_full, (NT - 1) & 1);
        tc_fence_after();
        const float inv = l_sum > 0.f ? 1.f / l_sum : 0.f;
        T* orow = reinterpret_cast<T*>(p.out) + static_cast<long long>(b) * p.o_bs + static_cast<long long>(qrow) * p.o_ls +
                  static_cast<long long>(h) * D;
        const uint32_t o_addr = tmem_base + lane_off + O_COL;
#pragma unroll 1
        for (int c = 0; c < DP; c += 16) {
            uint32_t raw[16];
            tmem_ld16(o_addr + c, raw);
            tmem_wait_ld();
            if (qrow < p.lq) {
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    if (c + 8 * g < D) {
                        float f8[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) f8[i] = __uint_as_float(raw[8 * g + i]) * inv;
                        *reinterpret_cast<uint4*>(orow + c + 8 * g) = pack8<T>(f8);
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<512>(tmem_base);
}

#include "attention_v2.cuh"
#include "attention_v3.cuh"

// Profiling aid: a device buffer of >= 4096 int64. CTA (0,0,0) of flash_attn3_kernel writes clock64() stamps for its
// first 32 KV tiles: [0..) MMA thread {P_A seen, A issued, P_B seen, B issued}, [1024..) softmax warp of tile A
// {S ready, exps done}, [2048..) the same for tile B. scripts/attn_timeline.py prints it.
extern "C" void ab_attn_set_debug(long long* buf) { ab::g_attn_dbg = buf; }
// AB_ATTN_V=2 selects the previous kernel (P through shared memory, two reads of S) for A/B measurements.
static int attn_version() {
    static const int v = [] { const char* e = getenv("AB_ATTN_V"); return e ? atoi(e) : 3; }();
    return v;
}
template <typename T, typename KV>
static int launch_attn_pp(const CUtensorMap& tq, const KV& kv, const AttnParams& p, cudaStream_t stream) {
    return attn_version() == 2 ? launch_attn2_multi<T, KV>(tq, kv, p, stream) : launch_attn3_multi<T, KV>(tq, kv, p, stream);
}
template <typename T>
static int launch_attn_pp1(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnParams& p,
                           cudaStream_t stream) {
    KvMaps<1> kv;
    kv.k[0] = tk;
    kv.v[0] = tv;
    return launch_attn_pp<T, KvMaps<1>>(tq, kv, p, stream);
}

template <int NCH, typename T>
static int launch_attn(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnParams& p,
                       cudaStream_t stream) {
    constexpr int STAGES = NCH >= 3 ? 1 : 2;
    constexpr int PBUF = NCH >= 3 ? 1 : 2;
    constexpr int smem = NCH * CHUNK_BYTES * (1 + 2 * STAGES) + PBUF * 2 * CHUNK_BYTES + 16 * 8 + 16 + 1024;
    if (ab_ensure_max_smem(reinterpret_cast<const void*>(flash_attn_kernel<NCH, T>), smem)) return -1;
    dim3 grid((p.lq + ATT_BM - 1) / ATT_BM, p.heads, p.batch);
    AB_CHECK_CUDA(launch_pdl(flash_attn_kernel<NCH, T>, grid, dim3(ATT_THREADS), smem, stream, tq, tk, tv, p));
    ab_count_launch(1);
    return 0;
}

}  // namespace ab

extern "C" int ab_flash_attn(const ab_attn_args* a, void* stream_) {
    using namespace ab;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(a && a->q && a->k && a->v && a->out, "null pointer");
    AB_REQUIRE(a->head_dim % 8 == 0 && a->head_dim >= 16 && a->head_dim <= 192, "head_dim must be a multiple of 8 in [16,192]");
    AB_REQUIRE(a->lq > 0 && a->lk > 0 && a->batch > 0 && a->heads > 0, "bad shape");
    AB_REQUIRE(a->q_ls % 8 == 0 && a->k_ls % 8 == 0 && a->v_ls % 8 == 0 && a->o_ls % 8 == 0, "row strides must be multiples of 8");
    AB_REQUIRE(a->q_bs % 8 == 0 && a->k_bs % 8 == 0 && a->v_bs % 8 == 0, "batch strides must be multiples of 8");
    const int D = a->head_dim;
    const int nch = (D + 63) / 64;
    AttnParams p{};
    p.dbg = g_attn_dbg;
    p.out = a->out; p.batch = a->batch; p.heads = a->heads; p.lq = a->lq; p.lk = a->lk; p.head_dim = D;
    p.o_bs = a->o_bs; p.o_ls = a->o_ls;
    const float scale = a->scale > 0.f ? a->scale : 1.0f / sqrtf(static_cast<float>(D));
    p.scale_log2 = scale * 1.4426950408889634f;
    p.causal = a->causal & 1;   // bit 1 of the field: choose the kernel as for a query length above 128 (see header)
    CUtensorMap tq, tk, tv;
    auto mk = [&](CUtensorMap* tm, const void* base, int L, long long bs, long long ls) {
        uint64_t dims[4] = {(uint64_t)D, (uint64_t)a->heads, (uint64_t)L, (uint64_t)a->batch};
        uint64_t str[3] = {(uint64_t)D * 2, (uint64_t)ls * 2, (uint64_t)(a->batch > 1 ? bs : ls * L) * 2};
        uint32_t box[4] = {64, 1, 128, 1};
        return ab_make_tensor_map(tm, a->dtype, 4, base, dims, str, box, 128);
    };
    if (mk(&tq, a->q, a->lq, a->q_bs, a->q_ls)) return -1;
    if (mk(&tk, a->k, a->lk, a->k_bs, a->k_ls)) return -1;
    if (mk(&tv, a->v, a->lk, a->v_bs, a->v_ls)) return -1;
#define AB_ATT(T)                                                      \
    (nch == 1 ? ((a->lq > 128 || (a->causal & 2)) ? launch_attn_pp1<T>(tq, tk, tv, p, stream) : launch_attn<1, T>(tq, tk, tv, p, stream)) \
     : nch == 2 ? launch_attn<2, T>(tq, tk, tv, p, stream)             \
                : launch_attn<3, T>(tq, tk, tv, p, stream))
    return a->dtype == AB_BF16 ? AB_ATT(__nv_bfloat16) : AB_ATT(__half);
#undef AB_ATT
}

// ---- patch-parallel attention: K/V spread over the GPUs of a subset, gathered by the kernel's own TMA loads ----
// k_ptrs[s] / v_ptrs[s] are the (local or peer-mapped) [B, lk_per_src, H, D] buffers of source s, all with the
// same strides; the global key order is source 0's keys, then source 1's, ...  See include/arbius_b200.h.
extern "C" int ab_flash_attn_multi(const ab_attn_args* a, const void* const* k_ptrs, const void* const* v_ptrs,
                                   int nsrc, int lk_per_src, void* stream_) {
    using namespace ab;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(a && a->q && a->out && k_ptrs && v_ptrs, "null pointer");
    AB_REQUIRE(nsrc == 1 || nsrc == 2 || nsrc == 4 || nsrc == 8, "nsrc must be 1, 2, 4 or 8");
    AB_REQUIRE(a->head_dim % 8 == 0 && a->head_dim >= 16 && a->head_dim <= 64, "multi-source attention needs head_dim in [16,64]");
    AB_REQUIRE(lk_per_src > 0 && lk_per_src % ATT_BN == 0, "keys per source must be a multiple of 128");
    AB_REQUIRE(!a->causal, "causal masking is not supported in the multi-source path");
    const int D = a->head_dim;
    AttnParams p{};
    p.dbg = g_attn_dbg;
    p.out = a->out; p.batch = a->batch; p.heads = a->heads; p.lq = a->lq; p.lk = nsrc * lk_per_src; p.head_dim = D;
    p.o_bs = a->o_bs; p.o_ls = a->o_ls;
    const float scale = a->scale > 0.f ? a->scale : 1.0f / sqrtf(static_cast<float>(D));
    p.scale_log2 = scale * 1.4426950408889634f;
    p.causal = 0;
    p.tiles_per_src = lk_per_src / ATT_BN;
    auto mk = [&](CUtensorMap* tm, const void* base, int L, long long bs, long long ls) {
        uint64_t dims[4] = {(uint64_t)D, (uint64_t)a->heads, (uint64_t)L, (uint64_t)a->batch};
        uint64_t str[3] = {(uint64_t)D * 2, (uint64_t)ls * 2, (uint64_t)(a->batch > 1 ? bs : ls * L) * 2};
        uint32_t box[4] = {64, 1, 128, 1};
        return ab_make_tensor_map(tm, a->dtype, 4, base, dims, str, box, 128);
    };
    CUtensorMap tq;
    if (mk(&tq, a->q, a->lq, a->q_bs, a->q_ls)) return -1;
#define AB_ATT_MULTI(T, N)                                                                   \
    {                                                                                         \
        KvMaps<N> kv;                                                                         \
        for (int s = 0; s < N; ++s) {                                                         \
            if (mk(&kv.k[s], k_ptrs[s], lk_per_src, a->k_bs, a->k_ls)) return -1;             \
            if (mk(&kv.v[s], v_ptrs[s], lk_per_src, a->v_bs, a->v_ls)) return -1;             \
        }                                                                                     \
        return launch_attn_pp<T, KvMaps<N>>(tq, kv, p, stream);                                   \
    }
#define AB_ATT_MULTI_T(T)                                  \
    switch (nsrc) {                                        \
        case 1: AB_ATT_MULTI(T, 1)                         \
        case 2: AB_ATT_MULTI(T, 2)                         \
        case 4: AB_ATT_MULTI(T, 4)                         \
        default: AB_ATT_MULTI(T, 8)                        \
    }
    if (a->dtype == AB_BF16) { AB_ATT_MULTI_T(__nv_bfloat16) } else { AB_ATT_MULTI_T(__half) }
#undef AB_ATT_MULTI_T
#undef AB_ATT_MULTI
}

// ---- segmented keys: [prefix | pixel bands on 1..8 GPUs] (see include/arbius_b200.h) ------------------------------
extern "C" int ab_flash_attn_segs(const ab_attn_segs_args* a, void* stream_) {
    using namespace ab;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(a && a->q && a->out, "null pointer");
    AB_REQUIRE(a->nsrc == 1 || a->nsrc == 2 || a->nsrc == 4 || a->nsrc == 8, "nsrc must be 1, 2, 4 or 8");
    AB_REQUIRE(a->head_dim % 8 == 0 && a->head_dim >= 16 && a->head_dim <= 64, "segmented attention needs head_dim in [16,64]");
    AB_REQUIRE(a->lq > 0 && a->batch > 0 && a->heads > 0, "bad shape");
    AB_REQUIRE(a->rows_per_src > 0 && (a->nsrc == 1 || a->rows_per_src % 8 == 0), "rows per source must be a multiple of 8");
    AB_REQUIRE(a->prefix_len >= 0 && (a->prefix_len == 0 || a->prefix), "prefix missing");
    AB_REQUIRE(a->q_ls % 8 == 0 && a->o_ls % 8 == 0 && a->pix_ls % 8 == 0 && a->pix_bs % 8 == 0, "strides must be multiples of 8");
    AB_REQUIRE(!a->wait_site || a->tick, "wait_site needs tick");
    const int D = a->head_dim;
    AttnParams p{};
    p.dbg = g_attn_dbg;
    p.out = a->out; p.batch = a->batch; p.heads = a->heads; p.lq = a->lq; p.head_dim = D;
    p.o_bs = a->o_bs; p.o_ls = a->o_ls;
    const float scale = a->scale > 0.f ? a->scale : 1.0f / sqrtf(static_cast<float>(D));
    p.scale_log2 = scale * 1.4426950408889634f;
    p.causal = 0;
    p.prefix_len = a->prefix_len; p.prefix_tiles = (a->prefix_len + ATT_BN - 1) / ATT_BN;
    p.pix_len = a->nsrc * a->rows_per_src; p.rows_per_src = a->rows_per_src; p.nsrc = a->nsrc;
    p.lk = p.prefix_len + p.pix_len;
    p.pk0 = a->prefix_k_head0; p.pv0 = a->prefix_v_head0; p.k0 = a->k_head0; p.v0 = a->v_head0;
    p.tick = a->tick; p.wait_site = a->wait_site; p.self_rank = a->self_rank;
    auto mk = [&](CUtensorMap* tm, const void* base, int heads_total, int L, long long bs, long long ls, uint32_t rows) {
        uint64_t dims[4] = {(uint64_t)D, (uint64_t)heads_total, (uint64_t)L, (uint64_t)a->batch};
        uint64_t str[3] = {(uint64_t)D * 2, (uint64_t)ls * 2, (uint64_t)(a->batch > 1 ? bs : ls * L) * 2};
        uint32_t box[4] = {64, 1, rows, 1};
        return ab_make_tensor_map(tm, a->dtype, 4, base, dims, str, box, 128);
    };
    CUtensorMap tq;
    if (mk(&tq, a->q, a->heads, a->lq, a->q_bs, a->q_ls, 128)) return -1;
    SegMaps kv;
    for (int s = 0; s < a->nsrc; ++s) {
        AB_REQUIRE(a->pix[s], "missing pixel source");
        if (mk(&kv.big[1 + s], a->pix[s], a->heads_total, a->rows_per_src, a->pix_bs, a->pix_ls, 128)) return -1;
        if (mk(&kv.small[1 + s], a->pix[s], a->heads_total, a->rows_per_src, a->pix_bs, a->pix_ls, 8)) return -1;
    }
    for (int s = a->nsrc; s < 8; ++s) { kv.big[1 + s] = kv.big[1]; kv.small[1 + s] = kv.small[1]; }
    if (a->prefix_len > 0) {
        if (mk(&kv.big[0], a->prefix, a->prefix_heads_total, a->prefix_len, a->prefix_bs, a->prefix_ls, 128)) return -1;
    } else {
        kv.big[0] = kv.big[1];
    }
    kv.small[0] = kv.small[1];
    if (a->dtype == AB_BF16) return launch_attn_pp<__nv_bfloat16, SegMaps>(tq, kv, p, stream);
    return launch_attn_pp<__half, SegMaps>(tq, kv, p, stream);
}
