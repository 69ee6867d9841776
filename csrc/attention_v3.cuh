// Flash attention v3 for head_dim <= 64: the v2 schedule (two 128-row query tiles per CTA, ping-ponged) with
//   * P kept in TENSOR MEMORY: the softmax warps write the packed 16-bit probabilities with tcgen05.st and P.V reads its
//     A operand from TMEM (tcgen05.mma [d], [a], b-desc). No swizzled STS, no proxy fence, and the P.V MMA fetches only V
//     from shared memory (the SS form needed 192 B/clk of shared memory for a 32-cycle MMA and ran at 48);
//   * ONE read of S per KV tile: the exponentials are taken against the running reference maximum of the previous tiles
//     while the same pass tracks the tile's maximum. Only when a row's maximum exceeds the reference by more than the
//     rescale threshold (a condition on that row's data alone) the warp repeats the tile in the two-pass way. A row
//     that did not jump computes the same bits in either way, so the result does not depend on its warp-mates.
// Included by attention.cu inside namespace ab, after attention_v2.cuh (helpers, KvMaps, SegMaps).
//
//   warp 0        TMA: Q_A, Q_B once; K_j/V_j ring (3 stages)
//   warp 1        MMA issuer: QK_A, QK_B, then per KV tile  PV_A, QK_A(next), PV_B, QK_B(next)
//   warps 2..5    softmax for tile A (TMEM lane quarters 2,3,0,1)
//   warps 6..9    softmax for tile B
// While one tile's softmax runs on the MUFU/FMA pipes the tensor core works on the other tile, and
// K/V shared-memory traffic is halved (both query tiles consume the same stage).
// TMEM: S_A [0,128) S_B [128,256) O_A [256,320) O_B [320,384) P_A [384,448) P_B [448,512) (two 16-bit P per column).


constexpr int A3_STAGES = 4;

template <typename T, typename KV>
__global__ void __launch_bounds__(A2_THREADS, 1)
flash_attn3_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ KV kv,
                   const AttnParams p) {
    constexpr int NSRC = KV::NSRC;
    constexpr float kRescaleThreshold = 8.0f;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;                                   // 2 x 16 KB
    uint8_t* sK = sQ + 2 * CHUNK_BYTES;                   // STAGES x 16 KB
    uint8_t* sV = sK + A3_STAGES * CHUNK_BYTES;           // STAGES x 16 KB
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + A3_STAGES * CHUNK_BYTES);
    uint64_t* q_full = bars;                  // 1
    uint64_t* kv_full = bars + 1;             // STAGES
    uint64_t* kv_empty = kv_full + A3_STAGES; // STAGES
    uint64_t* s_full = kv_empty + A3_STAGES;  // 2
    uint64_t* p_full = s_full + 2;            // 2
    uint64_t* o_full = p_full + 2;            // 2
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * 256;
    const int h = blockIdx.y;
    const int b = blockIdx.z;
    const int D = p.head_dim;
    const int DP = (D + 15) & ~15;
    const bool b_active = q0 + 128 < p.lq;
    long long* const dbg = (blockIdx.x | blockIdx.y | blockIdx.z) == 0 ? p.dbg : nullptr;
    int kv_len = p.lk;
    if (p.causal) kv_len = min(p.lk, q0 + 256);
    int NT = (kv_len + ATT_BN - 1) / ATT_BN;
    if constexpr (KV::kSegs) NT = p.prefix_tiles + (p.pix_len + ATT_BN - 1) / ATT_BN;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        if constexpr (KV::kSegs) {
            for (int s = 0; s <= p.nsrc; ++s) { tma_prefetch_desc(&kv.big[s]); tma_prefetch_desc(&kv.small[s]); }
        } else {
            for (int s = 0; s < NSRC; ++s) { tma_prefetch_desc(&kv.k[s]); tma_prefetch_desc(&kv.v[s]); }
        }
        mbar_init(q_full, 1);
        for (int s = 0; s < A3_STAGES; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
        for (int t = 0; t < 2; ++t) { mbar_init(&s_full[t], 1); mbar_init(&p_full[t], 128); mbar_init(&o_full[t], 1); }
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
        if (elect_one()) {
            mbar_expect_tx(q_full, (b_active ? 2 : 1) * CHUNK_BYTES);
            tma_load_4d(sQ, &tmQ, q_full, 0, h, q0, b);
            if (b_active) tma_load_4d(sQ + CHUNK_BYTES, &tmQ, q_full, 0, h, q0 + 128, b);
            int stage = 0; uint32_t phase = 0;
            unsigned int waited = 0;
            unsigned int tick = 0;
            if constexpr (KV::kSegs) { if (p.wait_site) tick = *p.tick; }
            for (int j = 0; j < NT; ++j) {
                mbar_wait(&kv_empty[stage], phase ^ 1);
                mbar_expect_tx(&kv_full[stage], 2 * CHUNK_BYTES);
                if constexpr (KV::kSegs) {
                    uint8_t* dk = sK + stage * CHUNK_BYTES;
                    uint8_t* dv = sV + stage * CHUNK_BYTES;
                    // a peer's K/V rows may only be read once its flag has arrived (one wait per source and CTA)
                    auto need = [&](int s) {
                        if (p.wait_site && s != p.self_rank && !((waited >> s) & 1u)) {
                            sync_wait_flags(p.wait_site, 1u << s, tick);
                            fence_proxy_async_all();
                            waited |= 1u << s;
                        }
                    };
                    if (j < p.prefix_tiles) {
                        tma_load_4d(dk, &kv.big[0], &kv_full[stage], 0, p.pk0 + h, j * ATT_BN, b);
                        tma_load_4d(dv, &kv.big[0], &kv_full[stage], 0, p.pv0 + h, j * ATT_BN, b);
                    } else {
                        const int g0 = (j - p.prefix_tiles) * ATT_BN;
                        const int g1 = min(g0 + ATT_BN, p.pix_len) - 1;           // last valid pixel row of the tile
                        const int s0 = g0 / p.rows_per_src, s1 = g1 / p.rows_per_src;
                        if (s0 == s1) {
                            // whole tile from one GPU (rows past the end of the last band are zero-filled by TMA)
                            need(s0);
                            tma_load_4d(dk, &kv.big[1 + s0], &kv_full[stage], 0, p.k0 + h, g0 - s0 * p.rows_per_src, b);
                            tma_load_4d(dv, &kv.big[1 + s0], &kv_full[stage], 0, p.v0 + h, g0 - s0 * p.rows_per_src, b);
                        } else {
                            // the tile straddles GPUs: sixteen 8-row boxes, each from the band that owns its rows
#pragma unroll 1
                            for (int bb = 0; bb < ATT_BN / 8; ++bb) {
                                const int g = g0 + 8 * bb;
                                int s = g / p.rows_per_src;
                                int row = g - s * p.rows_per_src;
                                if (g >= p.pix_len) { s = p.nsrc - 1; row = p.rows_per_src + 8; }   // out of bounds: zeros
                                else need(s);
                                tma_load_4d(dk + bb * 1024, &kv.small[1 + s], &kv_full[stage], 0, p.k0 + h, row, b);
                                tma_load_4d(dv + bb * 1024, &kv.small[1 + s], &kv_full[stage], 0, p.v0 + h, row, b);
                            }
                        }
                    }
                } else {
                    int src = 0, row = j * ATT_BN;
                    if (NSRC > 1) { src = j / p.tiles_per_src; row = (j - src * p.tiles_per_src) * ATT_BN; }
                    tma_load_4d(sK + stage * CHUNK_BYTES, &kv.k[src], &kv_full[stage], 0, h, row, b);
                    tma_load_4d(sV + stage * CHUNK_BYTES, &kv.v[src], &kv_full[stage], 0, h, row, b);
                }
                if (++stage == A3_STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // The whole warp runs this control flow and the elected lane issues inside `if (leader)`: descriptors and addresses
        // stay in uniform registers (under `if (lane == 0)` every UTCHMMA operand was moved out of a "divergent" context
        // with an ELECT / R2UR / BRA.U.ANY loop: 1280 cycles to issue one tile's twelve MMAs, which paced the kernel).
        // Measured alternatives (scripts/attn_timeline.py, 48x5x9216^2x64): the whole loop under one elected lane issues a
        // tile in 440 cycles instead of 740 but the two softmax groups then fall into closer lock-step (7.38 vs 6.96 ms);
        // software-pipelining the exponentials over chunks (7.54) and passing an explicit turn between the groups with
        // named barriers (8.54 exclusive, 7.52 half-way) did not beat it either.
        {
            const bool leader = elect_one();
            const uint32_t idesc_qk = make_idesc_f16(ATT_BM, ATT_BN, DT<T>::kBf16, false, false);
            const uint32_t idesc_pv = make_idesc_f16(ATT_BM, DP, DT<T>::kBf16, false, true);
            const uint32_t sq_base = smem_u32(sQ), sk_base = smem_u32(sK), sv_base = smem_u32(sV);
            auto issue_qk = [&](int t, int j) {
                const uint32_t q_addr = sq_base + t * CHUNK_BYTES;
                const uint32_t k_addr = sk_base + (j % A3_STAGES) * CHUNK_BYTES;
                if (leader) {
                    for (int ks = 0; ks < DP / 16; ++ks)
                        umma_ss(tmem_base + static_cast<uint32_t>(t) * 128u, make_smem_desc(q_addr + ks * 32, 16, 1024, kSwizzle128),
                                make_smem_desc(k_addr + ks * 32, 16, 1024, kSwizzle128), idesc_qk, ks > 0 ? 1u : 0u);
                    umma_commit(&s_full[t]);
                }
                __syncwarp();
            };
            auto wait_p = [&](int t, int j) {
                mbar_wait(&p_full[t], j & 1);
                tc_fence_after();
            };
            auto issue_pv = [&](int t, int j) {
                // A = P from tensor memory: 16 keys of a K step are 8 columns (two 16-bit values per column)
                const uint32_t p_taddr = tmem_base + 384u + static_cast<uint32_t>(t) * 64u;
                const uint32_t v_addr = sv_base + (j % A3_STAGES) * CHUNK_BYTES;
                if (leader) {
#pragma unroll
                    for (int ks = 0; ks < ATT_BN / 16; ++ks)
                        umma_ts(tmem_base + 256u + static_cast<uint32_t>(t) * 64u, p_taddr + ks * 8,
                                make_smem_desc(v_addr + ks * 16 * 128, CHUNK_BYTES, 1024, kSwizzle128), idesc_pv,
                                (j > 0 || ks > 0) ? 1u : 0u);
                    umma_commit(&o_full[t]);
                }
                __syncwarp();
            };
            mbar_wait(q_full, 0);
            mbar_wait(&kv_full[0], 0);
            tc_fence_after();
            issue_qk(0, 0);
            if (b_active) issue_qk(1, 0);
            for (int j = 0; j < NT; ++j) {
                const bool more = j + 1 < NT;
                if (more) {
                    mbar_wait(&kv_full[(j + 1) % A3_STAGES], ((j + 1) / A3_STAGES) & 1);
                    tc_fence_after();
                }
                // S_t is free as soon as P_t is written: the next QK goes first so the softmax warps get their next
                // scores one P.V earlier (they only need O again at their rescale point, after the max pass).
                wait_p(0, j);
                if (dbg && leader && j < 32) dbg[4 * j] = clock64();
                if (more) issue_qk(0, j + 1);
                issue_pv(0, j);
                if (dbg && leader && j < 32) dbg[4 * j + 1] = clock64();
                if (b_active) {
                    wait_p(1, j);
                    if (dbg && leader && j < 32) dbg[4 * j + 2] = clock64();
                    if (more) issue_qk(1, j + 1);
                    issue_pv(1, j);
                    if (dbg && leader && j < 32) dbg[4 * j + 3] = clock64();
                }
                // every MMA that read stage j has been issued; this commit fires when they are all done.
                // QK(j+1) issued above read stage j+1, which only delays the release slightly.
                if (leader) umma_commit(&kv_empty[j % A3_STAGES]);
                __syncwarp();
            }
        }
    } else {
        // ================= softmax warps: tile t, one thread per query row =================
        const int t = (warp - 2) >> 2;
        const int qd = warp & 3;
        const int r = qd * 32 + lane;
        const int qrow = q0 + t * 128 + r;
        const uint32_t lane_off = static_cast<uint32_t>(qd * 32) << 16;
        const bool active = (t == 0) || b_active;
        float m_ref = -INFINITY;
        float l_sum = 0.f;
        const uint32_t s_addr = tmem_base + lane_off + static_cast<uint32_t>(t) * 128u;
        const uint32_t o_addr = tmem_base + lane_off + 256u + static_cast<uint32_t>(t) * 64u;
        const uint32_t p_addr = tmem_base + lane_off + 384u + static_cast<uint32_t>(t) * 64u;
        if (active) {
            for (int j = 0; j < NT; ++j) {
                mbar_wait(&s_full[t], j & 1);
                tc_fence_after();
                if (dbg && r == 0 && j < 32) dbg[1024 * (t + 1) + 2 * j] = clock64();
                int col0 = j * ATT_BN;
                int col_hi = p.lk;
                if (p.causal) col_hi = min(col_hi, qrow + 1);
                if constexpr (KV::kSegs) {
                    // every tile counts its own valid keys: a short prefix tile, full pixel tiles, a short last one
                    col0 = 0;
                    col_hi = j < p.prefix_tiles ? min(ATT_BN, p.prefix_len - j * ATT_BN)
                                                : min(ATT_BN, p.pix_len - (j - p.prefix_tiles) * ATT_BN);
                }
                const bool full_tile = __all_sync(0xffffffffu, col0 + ATT_BN <= col_hi);
                uint32_t raw[2][32];
                bool o_ready = (j == 0);
                // P.V(j-1) has finished: O is stable and the P columns may be rewritten
                auto wait_o = [&]() {
                    if (!o_ready) {
                        mbar_wait(&o_full[t], (j - 1) & 1);
                        tc_fence_after();
                        o_ready = true;
                    }
                };
                // 32 probabilities -> 16 columns of packed 16-bit values in the P region
                auto store_p = [&](int c, const float (&pv)[32]) {
                    uint32_t pk[16];
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        float f8[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) f8[i] = pv[8 * g + i];
                        const uint4 u = pack8<T>(f8);
                        pk[4 * g] = u.x; pk[4 * g + 1] = u.y; pk[4 * g + 2] = u.z; pk[4 * g + 3] = u.w;
                    }
                    tmem_st16(p_addr + 16 * c, pk);
                };
                // warp-uniform: the first tile has no reference yet, a masked tile needs the column tests
                bool two_pass = (j == 0) || !full_tile;
                if (!two_pass) {
                    // ---- one pass against the reference of the previous tiles; the raw maximum rides along ----
                    const float neg_m = -m_ref;
                    float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
                    float2 sum01 = make_float2(0.f, 0.f), sum23 = make_float2(0.f, 0.f);
                    tmem_ld32(s_addr, raw[0]);
#pragma unroll
                    for (int c = 0; c < ATT_BN / 32; ++c) {
                        tmem_wait_ld();
                        if (c + 1 < ATT_BN / 32) tmem_ld32(s_addr + (c + 1) * 32, raw[(c + 1) & 1]);
                        float pv[32];
#pragma unroll
                        for (int i = 0; i < 32; i += 8) {
                            m0 = max3f(m0, __uint_as_float(raw[c & 1][i]), __uint_as_float(raw[c & 1][i + 1]));
                            m1 = max3f(m1, __uint_as_float(raw[c & 1][i + 2]), __uint_as_float(raw[c & 1][i + 3]));
                            m2 = max3f(m2, __uint_as_float(raw[c & 1][i + 4]), __uint_as_float(raw[c & 1][i + 5]));
                            m3 = max3f(m3, __uint_as_float(raw[c & 1][i + 6]), __uint_as_float(raw[c & 1][i + 7]));
                        }
#pragma unroll
                        for (int i = 0; i < 32; i += 2) {
                            const float2 t2 = ffma2(__uint_as_float(raw[c & 1][i]), __uint_as_float(raw[c & 1][i + 1]),
                                                    p.scale_log2, neg_m);
                            pv[i] = ex2_approx(t2.x);
                            pv[i + 1] = ex2_approx(t2.y);
                        }
#pragma unroll
                        for (int i = 0; i < 32; i += 4) {
                            sum01 = fadd2(sum01, pv[i], pv[i + 1]);
                            sum23 = fadd2(sum23, pv[i + 2], pv[i + 3]);
                        }
                        if (c == 0) wait_o();
                        store_p(c, pv);
                    }
                    const float mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
                    // the row jumped past the threshold (or has no reference yet): its exponentials above are void
                    const bool jump = (m_ref == -INFINITY) || (fmaf(mx, p.scale_log2, neg_m) > kRescaleThreshold);
                    if (__any_sync(0xffffffffu, jump)) {
                        two_pass = true;
                        tmem_wait_st();
                    } else {
                        l_sum += (sum01.x + sum01.y) + (sum23.x + sum23.y);
                    }
                }
                if (two_pass) {
                    // ---- pass 1: raw row max (chunk c+1 in flight while chunk c is reduced) ----
                    float mx = -INFINITY;
                    tmem_ld32(s_addr, raw[0]);
#pragma unroll
                    for (int c = 0; c < ATT_BN / 32; ++c) {
                        tmem_wait_ld();
                        if (c + 1 < ATT_BN / 32) tmem_ld32(s_addr + (c + 1) * 32, raw[(c + 1) & 1]);
                        else tmem_ld32(s_addr, raw[(c + 1) & 1]);               // chunk 0 again, for pass 2
                        if (full_tile) {
                            float m0 = mx, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
                            for (int i = 0; i < 32; i += 8) {
                                m0 = max3f(m0, __uint_as_float(raw[c & 1][i]), __uint_as_float(raw[c & 1][i + 1]));
                                m1 = max3f(m1, __uint_as_float(raw[c & 1][i + 2]), __uint_as_float(raw[c & 1][i + 3]));
                                m2 = max3f(m2, __uint_as_float(raw[c & 1][i + 4]), __uint_as_float(raw[c & 1][i + 5]));
                                m3 = max3f(m3, __uint_as_float(raw[c & 1][i + 6]), __uint_as_float(raw[c & 1][i + 7]));
                            }
                            mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
                        } else {
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                if (col0 + c * 32 + i < col_hi) mx = fmaxf(mx, __uint_as_float(raw[c & 1][i]));
                        }
                    }
                    // same test as the one-pass path (fma is monotonic, so max_i fma(s_i) == fma(max_i s_i))
                    bool need = false;
                    float factor = 1.f;
                    if (mx != -INFINITY &&
                        (m_ref == -INFINITY || fmaf(mx, p.scale_log2, -m_ref) > kRescaleThreshold)) {
                        const float m_new = mx * p.scale_log2;  // scale > 0: max commutes with the scaling
                        need = (m_ref != -INFINITY);
                        factor = need ? exp2f(m_ref - m_new) : 1.f;
                        m_ref = m_new;
                    }
                    const bool warp_need = __any_sync(0xffffffffu, need);
                    wait_o();
                    if (warp_need) {
                        l_sum *= factor;
#pragma unroll 1
                        for (int c = 0; c < DP; c += 16) {
                            uint32_t o16[16];
                            tmem_ld16(o_addr + c, o16);
                            tmem_wait_ld();
#pragma unroll
                            for (int i = 0; i < 16; ++i) o16[i] = __float_as_uint(__uint_as_float(o16[i]) * factor);
                            tmem_st16(o_addr + c, o16);
                        }
                    }
                    // ---- pass 2: p = 2^(s*c - m), row sum, P -> tensor memory ----
                    const float neg_m = (m_ref == -INFINITY) ? 0.f : -m_ref;
                    float2 sum01 = make_float2(0.f, 0.f), sum23 = make_float2(0.f, 0.f);
#pragma unroll
                    for (int c = 0; c < ATT_BN / 32; ++c) {
                        tmem_wait_ld();
                        if (c + 1 < ATT_BN / 32) tmem_ld32(s_addr + (c + 1) * 32, raw[(c + 1) & 1]);
                        float pv[32];
                        if (full_tile) {
#pragma unroll
                            for (int i = 0; i < 32; i += 2) {
                                const float2 t2 = ffma2(__uint_as_float(raw[c & 1][i]), __uint_as_float(raw[c & 1][i + 1]),
                                                        p.scale_log2, neg_m);
                                pv[i] = ex2_approx(t2.x);
                                pv[i + 1] = ex2_approx(t2.y);
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                pv[i] = (col0 + c * 32 + i < col_hi)
                                            ? ex2_approx(fmaf(__uint_as_float(raw[c & 1][i]), p.scale_log2, neg_m)) : 0.f;
                        }
#pragma unroll
                        for (int i = 0; i < 32; i += 4) {
                            sum01 = fadd2(sum01, pv[i], pv[i + 1]);
                            sum23 = fadd2(sum23, pv[i + 2], pv[i + 3]);
                        }
                        store_p(c, pv);
                    }
                    l_sum += (sum01.x + sum01.y) + (sum23.x + sum23.y);
                }
                tmem_wait_st();
                tc_fence_before();
                mbar_arrive(&p_full[t]);
                if (dbg && r == 0 && j < 32) dbg[1024 * (t + 1) + 2 * j + 1] = clock64();
            }
            // ---- epilogue ----
            mbar_wait(&o_full[t], (NT - 1) & 1);
            tc_fence_after();
            const float inv = l_sum > 0.f ? 1.f / l_sum : 0.f;
            T* orow = reinterpret_cast<T*>(p.out) + static_cast<long long>(b) * p.o_bs +
                      static_cast<long long>(qrow) * p.o_ls + static_cast<long long>(h) * D;
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
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<512>(tmem_base);
}

template <typename T, typename KV>
static int launch_attn3_multi(const CUtensorMap& tq, const KV& kv, const AttnParams& p, cudaStream_t stream) {
    constexpr int smem = (2 + 2 * A3_STAGES) * CHUNK_BYTES + 16 * 8 + 16 + 1024;
    if (ab_ensure_max_smem(reinterpret_cast<const void*>(flash_attn3_kernel<T, KV>), smem)) return -1;
    dim3 grid((p.lq + 255) / 256, p.heads, p.batch);
    AB_CHECK_CUDA(launch_pdl(flash_attn3_kernel<T, KV>, grid, dim3(A2_THREADS), smem, stream, tq, kv, p));
    ab_count_launch(1);
    return 0;
}

