// Lockstep persistent (min,+) Viterbi step kernel for sm_100a ("LS kernel") -- the batched
// "sequences as SpMM columns" step  v_t[i] = min_j fl(fl(E[o_t][i] + T'[i][j]) + v_{t-1}[j]).
//
// One persistent CTA per SM walks ALL time steps of Q = 4 or 8 sequences in lockstep:
//   * transitions (weight, source offset) are REGISTER resident for the whole kernel: a thread
//     owns up to 5 destination rows ("ranks") and keeps their in-edges in a statically indexed
//     register slice -- the model is read from memory exactly once per CTA;
//   * the score vectors live in shared memory, interleaved [state][4 sequences] and double
//     buffered, so one LDS.128 feeds four relaxations and the only per-relaxation memory
//     traffic is 4 bytes over the shared-memory crossbar.  K0 (sv_ls_plan.cpp) colours the
//     states so that the 8 lanes of every LDS.128 phase hit 8 distinct bank groups;
//   * the emission table E[symbol][state] lives in TENSOR MEMORY: TMEM lane = the thread's lane
//     in its warp quadrant, TMEM column = (warp group, symbol, rank).  The symbol of a sequence
//     is CTA-uniform, so E[o][own row] is one tcgen05.ld with a run-time column -- a datapath
//     that is separate from the shared-memory crossbar and costs no shared-memory capacity
//     (192 KB at N = 2405, which is what limited the ES kernel to one sequence per SM);
//   * one __syncthreads per time step (two with block-reduced heavy rows); lanes 0..Q-1 of
//     warp 0 are the sequence managers: they stream the symbols (8-byte loads, one step ahead)
//     and publish a small control block per step; finished sequences are written out and
//     replaced from a global longest-first queue, so ragged batches need no padding.
//
// Reference semantics: GraphBLAS_impl.cpp:59,64-73 (add order), GraphBLAS_helper.cpp:65
// (unreachable -> +inf).  Tensor cores are not used ((min,+) is not an MMA contraction); only
// their memory is.
#include "sv_ls_device.cuh"

namespace {

template <int TPG, int NQ4, typename SH, int NH, int HS>
__global__ void __launch_bounds__(TPG, 1)
sv_ls_kernel(SvLsDev m, SvSeqsDev s, float* __restrict__ out, float* __restrict__ best,
             int32_t* __restrict__ arg, unsigned* __restrict__ counter) {
    constexpr int Q = 4 * NQ4;
    constexpr int NW = TPG / 32;
    constexpr int MW = NW - 1; // manager warp: the one K0 gives the fewest slots
    constexpr int NL = SH::total();
    constexpr int NSL = NL + NH * HS;
    constexpr int RS = SV_LS_RSTRIDE;
    constexpr int NPAD = SV_ES_RMAX * TPG + 8; // positions per float4 group (static layout)
    constexpr int GS = NPAD * 16;              // bytes of one float4 group
    constexpr int BS = NQ4 * GS;               // bytes of one score buffer
    static_assert(SH::R == SV_ES_RMAX && SV_ES_RMAX == 5, "five ranks");
    static_assert(NH <= SV_ES_NHMAX && NH <= 2, "too many heavy rows");

    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* red = reinterpret_cast<float*>(smem_raw + 2 * BS); // [NH][NW][Q]
    __shared__ LsCtl ctl[4];
    __shared__ LsMgr mgr[8];
    __shared__ uint32_t s_taddr;
    __shared__ float s_bv[NW];
    __shared__ int s_bi[NW];

    const int tid = threadIdx.x, lane = tid & 31;
    // provably warp-uniform (uniform registers / uniform branches for everything derived from it)
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const unsigned long long* __restrict__ sym64 = static_cast<const unsigned long long*>(s.sym);
    const int nranks = m.ranks;
    const int n_pad = m.n_pad;

    // ---- one-time: tensor memory <- emission table, registers <- this thread's edges ----
    if (warp == 0) tmem_alloc(&s_taddr, static_cast<uint32_t>(m.tmem_cols));
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = s_taddr;
    // TMEM lanes of this warp's quadrant; columns of this warp group
    const uint32_t tE = tbase + (static_cast<uint32_t>(warp & 3) << 21) +
                        static_cast<uint32_t>(warp >> 2) * static_cast<uint32_t>(m.emits * RS);
    for (int o = 0; o < m.emits; ++o) {
        for (int r = 0; r < RS; ++r)
            tmem_st1(tE + static_cast<uint32_t>(o * RS + r),
                     r < nranks ? m.emis_p[static_cast<size_t>(o) * n_pad + r * TPG + tid] : 0.f);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");

    // emission scores of the heavy rows, indexed by the control block's column value (symbol * RS)
    constexpr int kHeStride = 256 * RS;
    __shared__ float s_he[NH > 0 ? NH * kHeStride : 1];
    const bool has_heavy = NH > 0 && m.n_heavy > 0;
    if constexpr (NH > 0) {
        for (int i = tid; i < NH * m.emits; i += TPG) {
            const int h = i / m.emits, o = i % m.emits;
            s_he[h * kHeStride + o * RS] = h < m.n_heavy ? m.emis_p[static_cast<size_t>(o) * n_pad + m.n_lpad + h] : 0.f;
        }
    }
    const uint32_t sbase = static_cast<uint32_t>(__cvta_generic_to_shared(smem_raw));
    float ew[NSL];
    uint32_t eoa[NSL];
#pragma unroll
    for (int k = 0; k < NSL; ++k) {
        ew[k] = m.edge_w[static_cast<size_t>(k) * TPG + tid];
        eoa[k] = m.edge_off[static_cast<size_t>(k) * TPG + tid] + sbase; // absolute address in buffer 0
    }
    auto eo = [&](int k) -> uint32_t { return eoa[k]; };
    uint32_t degs = 0; // 4 bits per rank
#pragma unroll
    for (int r = 0; r < SV_ES_RMAX; ++r) degs |= static_cast<uint32_t>(m.deg[warp * SV_ES_RMAX + r]) << (4 * r);
    const uint32_t own_s = sbase + static_cast<uint32_t>(tid) * 16u; // own row of rank 0, group 0, buffer 0

    // Control blocks are published TWO steps ahead (ring of 4): block tau+1 has been visible since
    // the barrier of step tau-1, so every warp can fetch its flags and the emission scores of the
    // next step before the barrier of step tau instead of after it.
    //
    // refill: hands new sequences to the manager lanes in `mask` and patches the control blocks of
    // the two coming steps d0, d1 (fresh: overwrite them).  Called by all lanes of the manager warp.
    unsigned long long mg_cur = 0, mg_nxt = 0;       // current / next 8 packed symbols
    const unsigned long long* mg_ptr = sym64;        // word that follows mg_nxt
    uint32_t mg_pos = 0, mg_t = 0, mg_len = 0;       // byte phase of the next symbol, last published local step, length (0 = idle)
    auto next_symbol = [&]() -> uint32_t {           // pure register arithmetic; the refreshing load is
        const uint32_t sym = static_cast<uint32_t>(mg_cur >> ((mg_pos & 7u) * 8u)) & 0xffu;
        if ((++mg_pos & 7u) == 0) {                  // consumed 8 steps after it is issued: never waited for
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
            // tail balance (SvSeqsDev::tail_begin): past that queue position the second float4 group stops fetching
            bool take = true;
            if (Q == 8 && lane >= 4) take = *reinterpret_cast<volatile unsigned*>(counter) < s.tail_begin;
            const unsigned w = take ? atomicAdd(counter, 1u) : 0xffffffffu;
            if (w < static_cast<unsigned>(s.nseq)) {
                sid = s.order[w];
                const uint64_t b = s.off[sid];
                mg_len = static_cast<uint32_t>(s.off[sid + 1] - b);
                mg_ptr = sym64 + (b >> 3);
                mg_cur = *mg_ptr++;
                mg_nxt = *mg_ptr++;
                mg_pos = static_cast<uint32_t>(b & 7u);
                col0 = next_symbol() * RS;
                started = true;
                one = mg_len == 1;
                if (mg_len > 1) {
                    col1 = next_symbol() * RS;
                    second = true;
                    two = mg_len == 2;
                    mg_t = 1;
                }
            }
            mgr[lane].sid = sid;
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
    if (warp == MW) refill((1u << Q) - 1u, ctl[0], ctl[1], true);
    __syncthreads();

    // Flags of the coming step and E[o_q][own rows] of its first group are fetched BEFORE the
    // barrier that ends the current step (here: before the loop).
    uint32_t et[NQ4][4][5];
    uint4 fl;       // x = init_mask, y = fin_mask, z = any_act of the coming step
    uint4 nfl, nc0; // ... in flight: flags and the first group's TMEM columns
    // Two stages, because under load a shared-memory read takes hundreds of cycles and the in-order
    // pipeline stalls at its first use: the control block is READ two small slices before the end
    // of the step, the tensor-memory fetch that needs it is ISSUED after the last slice.
    auto prefetch_ctl = [&](const LsCtl& nc) {
        nfl = *reinterpret_cast<const uint4*>(&nc.init_mask);
        nc0 = *reinterpret_cast<const uint4*>(&nc.col[0]);
    };
    auto prefetch_tmem = [&]() {
        fl = nfl;
        tmem_ld5(tE + nc0.x, et[0][0]);
        tmem_ld5(tE + nc0.y, et[0][1]);
        tmem_ld5(tE + nc0.z, et[0][2]);
        tmem_ld5(tE + nc0.w, et[0][3]);
    };
    auto prefetch_step = [&](const LsCtl& nc) {
        prefetch_ctl(nc);
        prefetch_tmem();
    };
    prefetch_step(ctl[0]);

    // Time loop.  tau = CTA step; par = tau & 1 = index of the score buffers that are read.
    uint32_t tau = 0;
    for (;;) {
        const uint32_t init_mask = fl.x, fin_mask = fl.y;
        if (fl.z == 0) break;
        const uint32_t par = tau & 1u;
        const uint32_t pofs = par * BS;
        unsigned char* vn = smem_raw + (par ^ 1u) * BS;
        // first gathers of the step (group 0: first heavy row if any, else rank 0) go out before anything else
        LsPair io;
        if (has_heavy)
            ls_gather<0>(eo, pofs, io, NH > 0 ? NL : 0, HS >= 2);
        else
            ls_gather<0>(eo, pofs, io, 0, SH::at(0) >= 2);
        const LsCtl& c = ctl[tau & 3u];
        [[maybe_unused]] uint4 cg1 = make_uint4(0u, 0u, 0u, 0u); // TMEM columns of group 1 (read in slice (0,3))

        // ---- managers first (off the tail of the step): control block of step tau + 2 ----
        if (warp == MW) {
            LsCtl& nx = ctl[(tau + 2u) & 3u];
            bool cont = false, last = false;
            uint32_t colv = 0;
            if (lane < Q && mg_t + 1u < mg_len) {
                cont = true;
                colv = next_symbol() * RS;
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

        // ---- rare: first step of a sequence, v_0 = E[o_0] + start overwrites the relaxed scores
        //      (same thread, program order) -- kept out of the hot path as a real branch ----
        auto init_fix = [&](int q0, int q1) {
#pragma unroll 1
            for (int r = 0; r < nranks; ++r) {
                const int p = r * TPG + tid;
                const float st = m.start_p[p];
#pragma unroll 1
                for (int q = q0; q < q1; ++q) {
                    if (!((init_mask >> q) & 1u)) continue;
                    const float ev = m.emis_p[static_cast<size_t>(c.col[q] / RS) * n_pad + p];
                    reinterpret_cast<float*>(vn + (q >> 2) * GS + p * 16)[q & 3] = ev + st;
                }
            }
        };

        // ---- light rows, one float4 group of 4 sequences at a time: slices (g, r) in order
        //      (0,0) (0,1) .. (0,4) (1,0) ..; E[o_q][own rows] of a group = 8 tensor-memory loads:
        //      group 0's were issued before the previous barrier, group 1's go out before the last
        //      slice of group 0 ----
        const uint32_t ownn = own_s + (pofs ^ BS); // own row, rank 0, group 0 of the buffer being written
        auto slice_step = [&](auto gc, auto rc) {
            constexpr int g = decltype(gc)::value;
            constexpr int r = decltype(rc)::value;
            constexpr bool last_r = r == SH::R - 1;
            if constexpr (r == 0) {
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 4; ++j) // ties the loaded registers to the wait above
                    asm volatile("" : "+r"(et[g][j][0]), "+r"(et[g][j][1]), "+r"(et[g][j][2]), "+r"(et[g][j][3]),
                                 "+r"(et[g][j][4]));
            }
            if constexpr (last_r && g + 1 < NQ4) {
                constexpr int g1 = g + 1 < NQ4 ? g + 1 : 0;
                tmem_ld5(tE + cg1.x, et[g1][0]);
                tmem_ld5(tE + cg1.y, et[g1][1]);
                tmem_ld5(tE + cg1.z, et[g1][2]);
                tmem_ld5(tE + cg1.w, et[g1][3]);
            }
            if constexpr (r == SH::R - 2) { // control reads, two small slices before they are needed
                if constexpr (g + 1 < NQ4)
                    cg1 = *reinterpret_cast<const uint4*>(&c.col[4 * (g + 1 < NQ4 ? g + 1 : 0)]);
                else
                    prefetch_ctl(ctl[(tau + 1u) & 3u]);
            }
            float e[4], acc[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) e[j] = __uint_as_float(et[g][j][r]);
            const int d = static_cast<int>((degs >> (4 * r)) & 15u);
            if constexpr (!last_r) { // next slice: next rank of the same group
                ls_run_slots<g * GS, SH::base(r), 0, SH::at(r), g * GS, SH::base(last_r ? 0 : r + 1),
                             (SH::at(last_r ? 0 : r + 1) >= 2), NSL>(d, ew, eo, pofs, pofs, e, acc, io);
            } else if constexpr (g + 1 < NQ4) { // next slice: rank 0 of group 1
                ls_run_slots<g * GS, SH::base(r), 0, SH::at(r), GS, 0, (SH::at(0) >= 2), NSL>(d, ew, eo, pofs, pofs, e, acc,
                                                                                             io);
            } else {
                ls_run_slots<g * GS, SH::base(r), 0, SH::at(r), 0, -1, false, NSL>(d, ew, eo, pofs, pofs, e, acc, io);
            }
            if (r < nranks) sts128<g * GS + r * TPG * 16>(ownn, acc);
        };
        auto group_step = [&](auto gc) {
            slice_step(gc, std::integral_constant<int, 0>{});
            slice_step(gc, std::integral_constant<int, 1>{});
            slice_step(gc, std::integral_constant<int, 2>{});
            slice_step(gc, std::integral_constant<int, 3>{});
            slice_step(gc, std::integral_constant<int, 4>{});
        };
        // ---- heavy rows (dense rows of profile HMMs): every thread relaxes a slice of the row and
        //      the warp minima go to shared memory FIRST in the step; after group 0 a barrier
        //      publishes them and warp h finishes row h while the other warps run group 1, so the
        //      finishing chain is off the critical path ----
        if constexpr (NH > 0) {
            if (has_heavy) {
                auto heavy_step = [&](auto kc) {
                    constexpr int k = decltype(kc)::value; // chain index: h-major over (h, g)
                    constexpr int h = k / NQ4, g = k % NQ4;
                    constexpr bool lastk = k + 1 == NH * NQ4;
                    constexpr int nh = lastk ? 0 : (k + 1) / NQ4, ng = lastk ? 0 : (k + 1) % NQ4;
                    float he[4], acc[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) he[j] = s_he[h * kHeStride + c.col[4 * g + j]];
                    const int d = h < m.n_heavy ? m.heavy_slots[h] : 0;
                    if constexpr (lastk) // next slice: light rows, group 0, rank 0
                        ls_run_slots<g * GS, NL + h * HS, 0, HS, 0, 0, (SH::at(0) >= 2), NSL>(d, ew, eo, pofs, pofs, he, acc, io);
                    else
                        ls_run_slots<g * GS, NL + h * HS, 0, HS, ng * GS, NL + nh * HS, (HS >= 2), NSL>(d, ew, eo, pofs, pofs, he,
                                                                                                  acc, io);
                    if (m.nonneg) { // one REDUX per value instead of five shuffle + min stages
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            acc[j] = __int_as_float(__reduce_min_sync(0xffffffffu, __float_as_int(acc[j])));
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
#pragma unroll
                            for (int dd = 16; dd > 0; dd >>= 1)
                                acc[j] = fminf(acc[j], __shfl_xor_sync(0xffffffffu, acc[j], dd));
                        }
                    }
                    if (lane == 0) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) red[(h * NW + warp) * Q + 4 * g + j] = acc[j];
                    }
                };
                heavy_step(std::integral_constant<int, 0>{});
                if constexpr (NH * NQ4 > 1) heavy_step(std::integral_constant<int, 1>{});
                if constexpr (NH * NQ4 > 2) heavy_step(std::integral_constant<int, 2>{});
                if constexpr (NH * NQ4 > 3) heavy_step(std::integral_constant<int, 3>{});
                static_assert(NH * NQ4 <= 4, "heavy chain");
            }
        }
        group_step(std::integral_constant<int, 0>{});
        if constexpr (NH > 0) {
            if (has_heavy) {
                __syncthreads();
                if (warp < m.n_heavy) {
                    // lane l: sequence q = l / LPQ, partials of warps (l % LPQ) + j * LPQ
                    constexpr int LPQ = 32 / Q;
                    const int h = warp, q = lane / LPQ, ph = m.n_lpad + h;
                    float x = pos_inf();
                    for (int w = lane % LPQ; w < NW; w += LPQ) x = fminf(x, red[(h * NW + w) * Q + q]);
#pragma unroll
                    for (int dd = LPQ / 2; dd > 0; dd >>= 1) x = fminf(x, __shfl_xor_sync(0xffffffffu, x, dd));
                    if ((init_mask >> q) & 1u)
                        x = m.emis_p[static_cast<size_t>(c.col[q] / RS) * n_pad + ph] + m.start_p[ph];
                    if (lane % LPQ == 0)
                        reinterpret_cast<float*>(vn + (q >> 2) * GS + ph * 16)[q & 3] = x;
                }
            }
        }
        if constexpr (NQ4 == 2) {
            // an idle second group -- the tail of a batch, SvSeqsDev::tail_begin -- is skipped (39.7 instead of 42.8 ms for
            // 4096 equal sequences on 148 SMs).  Its slices would have read the control block of the NEXT step; and the
            // emission fetch that group 0's last slice issued for it is retired with its destination registers still
            // allocated (an asynchronous tcgen05.ld must never land in registers the compiler has handed to something
            // else).  Guarding that fetch with a branch instead was measured: 42.4 ms, it disturbs the hot path.
            if ((fl.z & 0xf0u) != 0) {
                group_step(std::integral_constant<int, 1>{});
            } else {
                prefetch_ctl(ctl[(tau + 1u) & 3u]);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    asm volatile("" ::"r"(et[NQ4 - 1][j][0]), "r"(et[NQ4 - 1][j][1]), "r"(et[NQ4 - 1][j][2]),
                                 "r"(et[NQ4 - 1][j][3]), "r"(et[NQ4 - 1][j][4]));
            }
        }
        if (init_mask != 0) init_fix(0, Q);

        // first emission scores of step tau + 1 (control block published during step tau - 1, read in
        // slice (last group, 3)): in flight across the barrier
        prefetch_tmem();
        __syncthreads();

        // ---- rare path: write out finished sequences, hand out new ones ----
        if (fin_mask != 0) {
            for (int q = 0; q < Q; ++q) {
                if (!((fin_mask >> q) & 1u)) continue;
                const uint32_t sid = mgr[q].sid;
                const float* vq = reinterpret_cast<const float*>(vn + (q >> 2) * GS) + (q & 3);
                float bv = pos_inf();
                int bi = 0x7fffffff;
                for (int i = tid; i < m.n; i += TPG) {
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
                        s_bv[warp] = bv;
                        s_bi[warp] = bi;
                    }
                    __syncthreads();
                    if (warp == 0) {
                        bv = lane < NW ? s_bv[lane] : pos_inf();
                        bi = lane < NW ? s_bi[lane] : 0x7fffffff;
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
                    __syncthreads();
                }
            }
            __syncthreads(); // every thread is done with mgr[].sid
            if (warp == MW) refill(fin_mask, ctl[(tau + 1u) & 3u], ctl[(tau + 2u) & 3u], false);
            __syncthreads();
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); // retire the stale fetch, then redo it
            prefetch_step(ctl[(tau + 1u) & 3u]);
        }
        ++tau;
    }

    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); // the last prefetch is never consumed
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) tmem_free(tbase, static_cast<uint32_t>(m.tmem_cols));
}

} // namespace

// ------------------------------------------------------------- variant table ----
namespace {

using LsSparseT = LShape<8, 4, 3, 2, 1>; // random sparse models, tight staircase (chmm_gen k = 3)
using LsSparseM = LShape<8, 5, 4, 3, 2>;
using LsSparse = LShape<8, 6, 5, 4, 3>;  // random sparse models (in-degree capped at 8)
using LsProfile = LShape<4, 3, 3, 2, 2>; // <= 2 in-edges per light row + 2 dense rows (profile HMMs that do not
                                         // fit the chain kernel, sv_plain_chain.cu)

// order = order of preference in sv_ls_build_plan: fewer register slots first (the registers a
// tighter staircase frees hold gathers in flight).  Every staircase is compiled for 8 and for 4
// sequences per CTA; the launch picks one from the batch size.
const SvLsVariant kLsVariants[] = {
    // threads, shape, nh, hs, ranks
    {512, {8, 4, 3, 2, 1}, 0, 0, 5},
    {512, {8, 5, 4, 3, 2}, 0, 0, 5},
    {512, {8, 6, 5, 4, 3}, 0, 0, 5},
    {512, {4, 3, 3, 2, 2}, 2, 5, 5},
};
constexpr int kNumLsVariants = sizeof(kLsVariants) / sizeof(kLsVariants[0]);

const void* ls_variant_fn(int index, int nq4) {
    switch (index * 2 + (nq4 == 1 ? 1 : 0)) {
    case 0: return (const void*)sv_ls_kernel<512, 2, LsSparseT, 0, 0>;
    case 1: return (const void*)sv_ls_kernel<512, 1, LsSparseT, 0, 0>;
    case 2: return (const void*)sv_ls_kernel<512, 2, LsSparseM, 0, 0>;
    case 3: return (const void*)sv_ls_kernel<512, 1, LsSparseM, 0, 0>;
    case 4: return (const void*)sv_ls_kernel<512, 2, LsSparse, 0, 0>;
    case 5: return (const void*)sv_ls_kernel<512, 1, LsSparse, 0, 0>;
    case 6: return (const void*)sv_ls_kernel<512, 2, LsProfile, 2, 5>;
    case 7: return (const void*)sv_ls_kernel<512, 1, LsProfile, 2, 5>;
    default: return nullptr;
    }
}

} // namespace

int sv_ls_num_variants() { return kNumLsVariants; }
const SvLsVariant& sv_ls_variant(int index) { return kLsVariants[index]; }

int sv_ls_prepare(int index, int nq4, size_t smem) {
    const void* fn = ls_variant_fn(index, nq4);
    if (fn == nullptr) return cudaErrorInvalidValue;
    return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
}

int sv_ls_launch(int index, int nq4, const SvLsDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                 int32_t* d_arg, unsigned* d_counter, int grid, size_t smem, cudaStream_t st) {
    const void* fn = ls_variant_fn(index, nq4);
    if (fn == nullptr || s.sym_bytes != 1) return cudaErrorInvalidValue;
    void* args[] = {(void*)&m, (void*)&s, (void*)&d_out, (void*)&d_best, (void*)&d_arg, (void*)&d_counter};
    cudaError_t e = cudaLaunchKernel(fn, dim3(grid), dim3(kLsVariants[index].threads), args, smem, st);
    return e != cudaSuccess ? e : cudaGetLastError();
}
