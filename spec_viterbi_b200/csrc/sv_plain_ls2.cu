// Lockstep (min,+) Viterbi step kernel, TWO INDEPENDENT HALVES per CTA ("LS2") -- sm_100a.
//
// Same machinery as sv_plain_ls.cu (transitions in registers, score vectors as float4
// [position][4 sequences] in shared memory, emissions in tensor memory, software-pipelined
// gathers, register-resident sequence managers, control blocks two steps ahead), but the CTA
// is split into two halves of 256 threads that each own ALL rows (10 per thread) for their own
// 4 sequences and synchronise with their OWN named barrier.  The halves drift apart, so while
// one half drains its gather queue at its step barrier the other half is in the middle of a
// step: the shared-memory crossbar -- the resource that binds this kernel -- stays busy across
// step boundaries (with one thread set per CTA it idles ~20 % of the time around the barrier).
// Both halves read the same emission table in tensor memory (warp w and warp w + 8 share a
// TMEM quadrant and own the same rows).  The two source offsets of a slot pair share one
// register (16 bits each): 10 rows per thread need twice the slots of the 5-row kernels.
//
// Random sparse models only (no block-reduced heavy rows); K0 (sv_ls_plan.cpp) builds the
// same kind of plan for 256 threads x 10 ranks.
// Reference semantics: GraphBLAS_impl.cpp:59,64-73 (add order), GraphBLAS_helper.cpp:65.
#include "sv_ls_device.cuh"

namespace {

template <int TPG, typename SH>
__global__ void __launch_bounds__(2 * TPG, 1)
sv_ls2_kernel(SvLsDev m, SvSeqsDev s, float* __restrict__ out, float* __restrict__ best,
              int32_t* __restrict__ arg, unsigned* __restrict__ counter) {
    constexpr int Q = 4;                  // sequences per half
    constexpr int R = SH::R;              // rows per thread
    constexpr int NW = TPG / 32;          // warps per half
    constexpr int MW = NW - 1;            // manager warp of a half
    constexpr int NSL = SH::total();
    constexpr int NPAD = R * TPG + 8;     // positions per score buffer (static layout)
    constexpr int GS = NPAD * 16;         // bytes of one score buffer (float4 per position)
    constexpr int HB = 2 * GS;            // bytes of one half: two buffers
    static_assert(R == 10 && TPG == 256, "two halves of 256 threads, ten ranks");

    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ LsCtl ctl[2][4];
    __shared__ LsMgr mgr[2][Q];
    __shared__ uint32_t s_taddr;
    __shared__ float s_bv[2][NW];
    __shared__ int s_bi[2][NW];

    const int tid = threadIdx.x, lane = tid & 31, th = tid & (TPG - 1);
    // provably warp-uniform (uniform registers / uniform branches for everything derived from them)
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int half = warp / NW, wh = warp % NW;
    const unsigned long long* __restrict__ sym64 = static_cast<const unsigned long long*>(s.sym);
    const int nranks = m.ranks;
    const int n_pad = m.n_pad;
    auto half_sync = [&]() { // named barrier of this half (immediate ids: uniform branch)
        if (half == 0)
            asm volatile("bar.sync 1, %0;" ::"n"(TPG) : "memory");
        else
            asm volatile("bar.sync 2, %0;" ::"n"(TPG) : "memory");
    };

    // ---- one-time: tensor memory <- emission table (half 0), registers <- this thread's edges ----
    if (warp == 0) tmem_alloc(&s_taddr, static_cast<uint32_t>(m.tmem_cols));
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = s_taddr;
    // TMEM lanes of this warp's quadrant (warp w and w + 8 coincide); columns of this warp group
    const uint32_t tE = tbase + (static_cast<uint32_t>(wh & 3) << 21) +
                        static_cast<uint32_t>(wh >> 2) * static_cast<uint32_t>(m.emits * R);
    if (half == 0) {
        for (int o = 0; o < m.emits; ++o)
            for (int r = 0; r < R; ++r)
                tmem_st1(tE + static_cast<uint32_t>(o * R + r),
                         r < nranks ? m.emis_p[static_cast<size_t>(o) * n_pad + r * TPG + th] : 0.f);
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    // score buffers start as +inf
    for (int k = tid; k < 2 * HB / 16; k += 2 * TPG)
        reinterpret_cast<float4*>(smem_raw)[k] = make_float4(pos_inf(), pos_inf(), pos_inf(), pos_inf());

    const uint32_t hbase = static_cast<uint32_t>(__cvta_generic_to_shared(smem_raw)) + static_cast<uint32_t>(half) * HB;
    float ew[NSL];
    uint32_t eop[(NSL + 1) / 2]; // two 16-bit byte offsets (position * 16 < 65536) per register
#pragma unroll
    for (int k = 0; k < NSL; ++k) ew[k] = m.edge_w[static_cast<size_t>(k) * TPG + th];
#pragma unroll
    for (int k = 0; k < NSL; k += 2) {
        const uint32_t lo = m.edge_off[static_cast<size_t>(k) * TPG + th];
        const uint32_t hi = k + 1 < NSL ? m.edge_off[static_cast<size_t>(k + 1) * TPG + th] : 0u;
        eop[k >> 1] = lo | (hi << 16);
    }
    auto eo = [&](int k) -> uint32_t { return (k & 1) ? (eop[k >> 1] >> 16) : (eop[k >> 1] & 0xffffu); };
    uint32_t degs_lo = 0, degs_hi = 0; // 4 bits per rank: ranks 0..7, 8..9
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const uint32_t d = m.deg[wh * R + r];
        if (r < 8)
            degs_lo |= d << (4 * r);
        else
            degs_hi |= d << (4 * (r - 8));
    }
    const uint32_t own_s = hbase + static_cast<uint32_t>(th) * 16u; // own row of rank 0, buffer 0

    // ---- sequence managers: lanes 0..3 of the last warp of each half, state in registers ----
    unsigned long long mg_cur = 0, mg_nxt = 0;
    const unsigned long long* mg_ptr = sym64;
    uint32_t mg_pos = 0, mg_t = 0, mg_len = 0;
    auto next_symbol = [&]() -> uint32_t {
        const uint32_t sym = static_cast<uint32_t>(mg_cur >> ((mg_pos & 7u) * 8u)) & 0xffu;
        if ((++mg_pos & 7u) == 0) { // consumed 8 steps after it is issued: never waited for
            mg_cur = mg_nxt;
            mg_nxt = *mg_ptr++;
        }
        return sym;
    };
    auto refill = [&](uint32_t mask, LsCtl& d0, LsCtl& d1, bool fresh) {
        bool started = false, one = false, second = false, two = false;
        uint32_t col0 = 0, col1 = 0;
        if (lane < Q && ((mask >> lane) & 1u)) {
            mg_len = 0;
            mg_t = 0;
            uint32_t sid = 0;
            const unsigned w = atomicAdd(counter, 1u);
            if (w < static_cast<unsigned>(s.nseq)) {
                sid = s.order[w];
                const uint64_t b = s.off[sid];
                mg_len = static_cast<uint32_t>(s.off[sid + 1] - b);
                mg_ptr = sym64 + (b >> 3);
                mg_cur = *mg_ptr++;
                mg_nxt = *mg_ptr++;
                mg_pos = static_cast<uint32_t>(b & 7u);
                col0 = next_symbol() * R;
                started = true;
                one = mg_len == 1;
                if (mg_len > 1) {
                    col1 = next_symbol() * R;
                    second = true;
                    two = mg_len == 2;
                    mg_t = 1;
                }
            }
            mgr[half][lane].sid = sid;
        }
        const uint32_t startm = __ballot_sync(0xffffffffu, started);
        const uint32_t onem = __ballot_sync(0xffffffffu, one);
        const uint32_t secm = __ballot_sync(0xffffffffu, second);
        const uint32_t twom = __ballot_sync(0xffffffffu, two);
        if (lane < Q && (fresh || ((mask >> lane) & 1u))) {
            d0.col[lane] = col0;
            d1.col[lane] = col1;
        }
        if (lane == 0) {
            if (fresh) {
                d0.init_mask = startm; d0.fin_mask = onem; d0.any_act = startm;
                d1.init_mask = 0;      d1.fin_mask = twom; d1.any_act = secm;
            } else {
                d0.init_mask |= startm; d0.fin_mask |= onem; d0.any_act |= startm;
                d1.fin_mask |= twom;    d1.any_act |= secm;
            }
        }
        __syncwarp();
    };
    if (wh == MW) refill((1u << Q) - 1u, ctl[half][0], ctl[half][1], true);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads(); // emission table, +inf buffers and the first control blocks are visible to both halves
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // E[o_q][own rows]: ranks 0..4 (eta) are fetched before the barrier that ends the previous step,
    // ranks 5..9 (etb) while slice 4 runs
    uint32_t eta[4][5], etb[4][5];
    uint4 fl;       // x = init_mask, y = fin_mask, z = any_act of the coming step
    uint4 nfl, nc0; // ... in flight
    auto prefetch_ctl = [&](const LsCtl& nc) {
        nfl = *reinterpret_cast<const uint4*>(&nc.init_mask);
        nc0 = *reinterpret_cast<const uint4*>(&nc.col[0]);
    };
    auto prefetch_tmem = [&]() {
        fl = nfl;
        tmem_ld5(tE + nc0.x, eta[0]);
        tmem_ld5(tE + nc0.y, eta[1]);
        tmem_ld5(tE + nc0.z, eta[2]);
        tmem_ld5(tE + nc0.w, eta[3]);
    };
    prefetch_ctl(ctl[half][0]);
    prefetch_tmem();

    // Time loop of this half.  tau = step of the half; par = tau & 1 = score buffer that is read.
    uint32_t tau = 0;
    for (;;) {
        const uint32_t init_mask = fl.x, fin_mask = fl.y;
        if (fl.z == 0) break;
        const uint32_t par = tau & 1u;
        const uint32_t pofs = hbase + par * GS; // absolute shared address of the buffer that is read
        unsigned char* vn = smem_raw + half * HB + (par ^ 1u) * GS;
        LsPair io; // first gathers of the step go out before anything else
        ls_gather<0>(eo, pofs, io, 0, SH::at(0) >= 2);
        const LsCtl& c = ctl[half][tau & 3u];

        // ---- managers first (off the tail of the step): control block of step tau + 2 ----
        if (wh == MW) {
            LsCtl& nx = ctl[half][(tau + 2u) & 3u];
            bool cont = false, last = false;
            uint32_t colv = 0;
            if (lane < Q && mg_t + 1u < mg_len) {
                cont = true;
                colv = next_symbol() * R;
                last = mg_t + 2u == mg_len;
                ++mg_t;
            }
            const uint32_t contm = __ballot_sync(0xffffffffu, cont);
            const uint32_t lastm = __ballot_sync(0xffffffffu, last);
            if (lane < Q) nx.col[lane] = colv;
            if (lane == 0) {
                nx.init_mask = 0;
                nx.fin_mask = lastm;
                nx.any_act = contm;
            }
        }

        const uint32_t ownn = own_s + (par ^ 1u) * GS; // own row of rank 0 in the buffer being written
        auto slice_step = [&](auto rc) {
            constexpr int r = decltype(rc)::value;
            constexpr bool last_r = r == R - 1;
            if constexpr (r == 0) {
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 4; ++j) // ties the loaded registers to the wait above
                    asm volatile("" : "+r"(eta[j][0]), "+r"(eta[j][1]), "+r"(eta[j][2]), "+r"(eta[j][3]), "+r"(eta[j][4]));
            }
            if constexpr (r == 4) { // emission scores of ranks 5..9 (nc0 still holds this step's columns)
                tmem_ld5(tE + nc0.x + 5u, etb[0]);
                tmem_ld5(tE + nc0.y + 5u, etb[1]);
                tmem_ld5(tE + nc0.z + 5u, etb[2]);
                tmem_ld5(tE + nc0.w + 5u, etb[3]);
            }
            if constexpr (r == 5) {
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    asm volatile("" : "+r"(etb[j][0]), "+r"(etb[j][1]), "+r"(etb[j][2]), "+r"(etb[j][3]), "+r"(etb[j][4]));
            }
            if constexpr (r == R - 2) prefetch_ctl(ctl[half][(tau + 1u) & 3u]); // two small slices early
            float e[4], acc[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) e[j] = __uint_as_float(r < 5 ? eta[j][r < 5 ? r : 0] : etb[j][r < 5 ? 0 : r - 5]);
            const int d = static_cast<int>(r < 8 ? (degs_lo >> (4 * (r & 7))) & 15u : (degs_hi >> (4 * (r & 1))) & 15u);
            if constexpr (!last_r)
                ls_run_slots<0, SH::base(r), 0, SH::at(r), 0, SH::base(last_r ? 0 : r + 1), (SH::at(last_r ? 0 : r + 1) >= 2), NSL>(
                    d, ew, eo, pofs, pofs, e, acc, io);
            else
                ls_run_slots<0, SH::base(r), 0, SH::at(r), 0, -1, false, NSL>(d, ew, eo, pofs, pofs, e, acc, io);
            if (r < nranks) sts128<r * TPG * 16>(ownn, acc);
        };
        slice_step(std::integral_constant<int, 0>{});
        slice_step(std::integral_constant<int, 1>{});
        slice_step(std::integral_constant<int, 2>{});
        slice_step(std::integral_constant<int, 3>{});
        slice_step(std::integral_constant<int, 4>{});
        slice_step(std::integral_constant<int, 5>{});
        slice_step(std::integral_constant<int, 6>{});
        slice_step(std::integral_constant<int, 7>{});
        slice_step(std::integral_constant<int, 8>{});
        slice_step(std::integral_constant<int, 9>{});

        // ---- rare: first step of a sequence, v_0 = E[o_0] + start overwrites the relaxed scores
        //      (same thread, program order) ----
        if (init_mask != 0) {
#pragma unroll 1
            for (int r = 0; r < nranks; ++r) {
                const int p = r * TPG + th;
                const float st = m.start_p[p];
#pragma unroll 1
                for (int q = 0; q < Q; ++q) {
                    if (!((init_mask >> q) & 1u)) continue;
                    const float ev = m.emis_p[static_cast<size_t>(c.col[q] / R) * n_pad + p];
                    reinterpret_cast<float*>(vn + p * 16)[q] = ev + st;
                }
            }
        }

        // first emission scores of step tau + 1: in flight across the barrier
        prefetch_tmem();
        half_sync();

        // ---- rare path: write out finished sequences, hand out new ones ----
        if (fin_mask != 0) {
            for (int q = 0; q < Q; ++q) {
                if (!((fin_mask >> q) & 1u)) continue;
                const uint32_t sid = mgr[half][q].sid;
                const float* vq = reinterpret_cast<const float*>(vn) + q;
                float bv = pos_inf();
                int bi = 0x7fffffff;
                for (int i = th; i < m.n; i += TPG) {
                    float x = pos_inf();
                    const uint32_t c1 = m.copy_ptr[i + 1];
                    for (uint32_t cc = m.copy_ptr[i]; cc < c1; ++cc) x = fminf(x, vq[m.copy_p[cc] * 4u]);
                    if (out != nullptr) out[static_cast<size_t>(sid) * m.n + i] = x;
                    if (x < bv) {
                        bv = x;
                        bi = i;
                    }
                }
                if (best != nullptr || arg != nullptr) {
#pragma unroll
                    for (int d = 16; d > 0; d >>= 1) {
                        const float ov = __shfl_xor_sync(0xffffffffu, bv, d);
                        const int oi = __shfl_xor_sync(0xffffffffu, bi, d);
                        if (ov < bv || (ov == bv && oi < bi)) {
                            bv = ov;
                            bi = oi;
                        }
                    }
                    if (lane == 0) {
                        s_bv[half][wh] = bv;
                        s_bi[half][wh] = bi;
                    }
                    half_sync();
                    if (wh == 0) {
                        bv = lane < NW ? s_bv[half][lane] : pos_inf();
                        bi = lane < NW ? s_bi[half][lane] : 0x7fffffff;
#pragma unroll
                        for (int d = 16; d > 0; d >>= 1) {
                            const float ov = __shfl_xor_sync(0xffffffffu, bv, d);
                            const int oi = __shfl_xor_sync(0xffffffffu, bi, d);
                            if (ov < bv || (ov == bv && oi < bi)) {
                                bv = ov;
                                bi = oi;
                            }
                        }
                        if (lane == 0) {
                            if (best != nullptr) best[sid] = bv;
                            if (arg != nullptr) arg[sid] = bi == 0x7fffffff ? -1 : bi;
                        }
                    }
                    half_sync();
                }
            }
            half_sync(); // every thread of the half is done with mgr[half][].sid
            if (wh == MW) refill(fin_mask, ctl[half][(tau + 1u) & 3u], ctl[half][(tau + 2u) & 3u], false);
            half_sync();
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); // retire the stale fetch, then redo it
            prefetch_ctl(ctl[half][(tau + 1u) & 3u]);
            prefetch_tmem();
        }
        ++tau;
    }

    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); // the last prefetch is never consumed
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads(); // both halves are done
    if (warp == 0) tmem_free(tbase, static_cast<uint32_t>(m.tmem_cols));
}

using Ls2Sparse = LShape<8, 5, 4, 4, 3, 3, 2, 2, 1, 1>;

} // namespace

const void* sv_ls2_variant_fn(int which) {
    switch (which) {
    case 0: return (const void*)sv_ls2_kernel<256, Ls2Sparse>;
    default: return nullptr;
    }
}
