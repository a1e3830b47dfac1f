// Chain kernel for sm_100a: the plain (min,+) Viterbi step for "chain + specials" models (every
// profile HMM the reference ships), with the state vectors of Q sequences held in REGISTERS.
//
//   v_t[i] = min_j fl(fl(E[o_t][i] + T'[i][j]) + v_{t-1}[j])          (GraphBLAS_impl.cpp:59,64-73)
//
// One persistent CTA walks all time steps of Q = 8 / 4 / 2 / 1 sequences in lockstep:
//   * a thread owns SV_CHAIN_RPT = 5 consecutive chain positions; v_{t-1} of its positions for all Q
//     sequences lives in registers and is updated in place, so the band edge (p-1) -> p is a register
//     read; only position 0 of a thread needs its neighbour's last score, which every thread publishes
//     as one 16-byte shared-memory vector per 4 sequences and step (read back the same way);
//   * the hub state's score (special 0, the N state) is a broadcast shared-memory read;
//   * the two dense rows (N and C) need only min_k v[k] over the chain because all their chain
//     in-edges carry one weight (K0 checks it): fl(c + x) is monotone, so
//     min_k fl(fl(E+u) + v[k]) == fl(fl(E+u) + min_k v[k]) bit for bit.  Each thread folds its five
//     new scores into a running minimum, one integer REDUX per sequence reduces it over the warp
//     (scores >= 0: float order = signed-int order; else through an order-preserving key), warp w < Q finishes sequence w's two
//     special rows in the NEXT step from the per-warp partials -- one barrier per step;
//   * the inner add of fl(fl(E + T') + v) does not depend on the sequence, so A = fl(E[o][p] + w_pred[p]) and
//     B = fl(E[o][p] + w_hub[p]) are TABULATED once per CTA (same IEEE add, same bits): A and the first B of a thread
//     live in TENSOR MEMORY (lane = thread's lane in its warp quadrant, column = (warp group, symbol, 6 entries)),
//     fetched as one tcgen05.ld.x4 + .x2 with a run-time column one sequence ahead of its use; the other four B live
//     in shared memory as one float4 per (symbol, thread), fetched two sequences ahead (the two tables are 385 KB for
//     2405.chmm; tensor memory holds 256 KB);
//   * sequence managers (lane q of the last warp) stream the symbols and publish a control block two
//     steps ahead, exactly as in the lockstep kernel (sv_plain_ls.cu); finished sequences are
//     written out with their (best, argmin) and replaced from a global longest-first queue.
//
// Per (position, sequence) and step: 2 IEEE adds (8 of 10 per sequence as packed FADD2) + 1 FMNMX + 0.4
// FMNMX3 for FOUR transitions -- no per-relaxation shared-memory traffic is left, so the kernel is bound by
// FP32 issue.
// Tensor cores are not used ((min,+) is not an MMA contraction); only their memory is.
#include "sv_ls_device.cuh"

#include <cstdlib>

namespace {

constexpr int kChainHeStride = 256; // special-row emissions indexed by symbol
// Tensor-memory columns per symbol (see "tabulated level-1 entries" below): A of positions 1,2,3,4,0 and B of position 0
constexpr int kChainCS = SV_CHAIN_TMEM_CS;
static_assert(kChainCS == SV_CHAIN_RPT + 1, "layout");

// N consecutive floats of shared memory as ONE load / store (N = 1, 2, 4; p aligned to 4 N bytes)
template <int N> __device__ __forceinline__ void lds_vec(const float* p, float (&x)[N]) {
    if constexpr (N == 4) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        x[0] = t.x; x[1] = t.y; x[2] = t.z; x[3] = t.w;
    } else if constexpr (N == 2) {
        const float2 t = *reinterpret_cast<const float2*>(p);
        x[0] = t.x; x[1] = t.y;
    } else {
        x[0] = *p;
    }
}
template <int N> __device__ __forceinline__ void sts_vec(float* p, const float (&x)[N]) {
    if constexpr (N == 4) *reinterpret_cast<float4*>(p) = make_float4(x[0], x[1], x[2], x[3]);
    else if constexpr (N == 2) *reinterpret_cast<float2*>(p) = make_float2(x[0], x[1]);
    else *p = x[0];
}

// Warp-wide minimum on the integer REDUX unit.  NN (every score >= 0 or -0 / +inf, checked by K0): the signed-int
// order of the bit patterns IS the float order.  Otherwise (two of the reference's shipped models, 200.chmm and
// 400.chmm, carry one transition with "probability" > 1, i.e. a negative weight) the bits are first mapped to an
// order-preserving key -- flip the low 31 bits of negative values -- and mapped back (the map is an involution).
template <bool NN> __device__ __forceinline__ float warp_min_redux(float x) {
    int i = __float_as_int(x);
    if constexpr (!NN) i ^= (i >> 31) & 0x7fffffff;
    i = __reduce_min_sync(0xffffffffu, i);
    if constexpr (!NN) i ^= (i >> 31) & 0x7fffffff;
    return __int_as_float(i);
}

// Packed IEEE adds (sm_100+: add.rn.f32x2 -> FADD2): two independent round-to-nearest fp32 adds in ONE
// instruction on a 64-bit register pair.  Each half is an ordinary IEEE add, so the bits are the oracle's; 16 of
// the 20 adds per sequence pair up (positions 1,2 and 3,4 of a thread with their band predecessors 0,1 and 2,3).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 p, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p));
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

// E of the thread's five positions for one sequence as ONE tcgen05.ld.x8 (three columns more than needed: they
// belong to the next symbol, or to the unused tail of the allocation, and are ignored)
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
}

// A of the thread's five positions and B of its first one for one sequence: 6 consecutive TMEM columns
__device__ __forceinline__ void tmem_ld6(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(taddr)
                 : "memory");
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0,%1}, [%2];" : "=r"(r[4]), "=r"(r[5]) : "r"(taddr + 4u) : "memory");
}

template <int TPG, int Q, bool NN, bool PK>
__global__ void __launch_bounds__(TPG, 512 / TPG)
sv_chain_kernel(SvChainDev m, SvSeqsDev s, float* __restrict__ out, float* __restrict__ best,
                int32_t* __restrict__ arg, unsigned* __restrict__ counter) {
    constexpr int NW = TPG / 32;
    constexpr int MW = NW - 1; // manager warp (its chain positions are mostly padding)
    constexpr int R = SV_CHAIN_RPT;
    constexpr int CS = kChainCS;
    static_assert(PK, "the tabulated layout is the packed one");
    constexpr int QS = 8; // stride of the per-sequence shared arrays
    // one tcgen05.ld.x8 per sequence (5 % off the single-sequence step) where registers allow, else .x4 + .x1
    constexpr bool X8 = Q <= 2;
    static_assert(R == 5 && Q <= 8 && NW <= 32, "layout");

    __shared__ LsCtl ctl[4];
    __shared__ LsMgr mgr[8];
    __shared__ uint32_t s_taddr;
    constexpr int QV = Q >= 4 ? 4 : Q; // sequences per shared-memory vector (float4 / float2 / float)
    constexpr int NG = Q / QV;         // vectors per thread and step
    __shared__ __align__(16) float s_red[2][NW][QS]; // [parity][warp][q] chain minimum of the warp
    // s_top[parity][group][1 + tid] = score of thread tid's LAST position for the QV sequences of the group: the
    // band edge into the next thread's first position.  Entry 0 (the predecessor of position 0) stays +inf.
    __shared__ __align__(16) float s_top[2][NG][(TPG + 1) * QV];
    __shared__ __align__(16) float s_vs[2][SV_CHAIN_NSP][QS]; // [parity][special][q]
    __shared__ float s_he[SV_CHAIN_NSP * kChainHeStride];
    // B = fl(E + w_hub) of positions 1..4 of every thread, per symbol: [symbol][thread] float4 (dynamic shared memory)
    extern __shared__ __align__(16) unsigned char chain_dyn_smem[];
    float4* const sB = reinterpret_cast<float4*>(chain_dyn_smem);
    __shared__ float s_bv[NW];
    __shared__ int s_bi[NW];

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0); // provably warp-uniform
    const unsigned long long* __restrict__ sym64 = static_cast<const unsigned long long*>(s.sym);
    const int n_pos = m.n_pos;

    // ---- one-time: tensor memory <- emission table; registers <- this thread's weights ----
    if (warp == 0) tmem_alloc(&s_taddr, static_cast<uint32_t>(m.tmem_cols));
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = s_taddr;
    const uint32_t tE = tbase + (static_cast<uint32_t>(warp & 3) << 21) +
                        static_cast<uint32_t>(warp >> 2) * static_cast<uint32_t>(m.emits * CS);
    // Tabulated level-1 entries: the step needs E only inside A = fl(E[o][p] + w_pred[p]) and B = fl(E[o][p] + w_hub[p])
    // (GraphBLAS_impl.cpp:65,69: the inner add of fl(fl(E + T') + v)), which do not depend on the sequence -- so they are
    // formed ONCE here, with the same IEEE add the step used to issue per sequence and step, and the step is left with
    // one add per transition.  Both tables are 4 N emits bytes (192 KB each for 2405.chmm): A lives in tensor memory
    // (column order of a symbol: positions 1,2,3,4,0, so that the x4 load delivers the pairs (1,2) and (3,4) in aligned
    // register pairs) together with B of position 0; B of positions 1..4 lives in shared memory as one float4 per
    // (symbol, thread) -- tensor memory holds 256 KB, the two tables need 385 KB.
    float wp[R], wh[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        wp[r] = m.w_pred[tid * R + r];
        wh[r] = m.w_hub[tid * R + r];
    }
    for (int o = 0; o < m.emits; ++o) {
        float e[R];
#pragma unroll
        for (int r = 0; r < R; ++r) e[r] = m.emis_p[static_cast<size_t>(o) * n_pos + tid * R + r];
#pragma unroll
        for (int r = 0; r < R; ++r) tmem_st1(tE + static_cast<uint32_t>(o * CS + (r == 0 ? R - 1 : r - 1)), e[r] + wp[r]);
        tmem_st1(tE + static_cast<uint32_t>(o * CS + R), e[0] + wh[0]);
        sB[o * TPG + tid] = make_float4(e[1] + wh[1], e[2] + wh[2], e[3] + wh[3], e[4] + wh[4]);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    for (int i = tid; i < SV_CHAIN_NSP * m.emits; i += TPG) {
        const int h = i / m.emits, o = i % m.emits;
        s_he[h * kChainHeStride + o] = m.emis_s[i];
    }
    for (int i = tid; i < 2 * NW * QS; i += TPG) (&s_red[0][0][0])[i] = pos_inf();
    for (int i = tid; i < 2 * NG * (TPG + 1) * QV; i += TPG) (&s_top[0][0][0])[i] = pos_inf();
    if (tid < 2 * SV_CHAIN_NSP * QS) (&s_vs[0][0][0])[tid] = pos_inf();

    // special rows: lane s of a warp finishes special s
    // (no run-time indexing into the kernel parameters: that would move the whole struct to local memory)
    static_assert(SV_CHAIN_NSP == 2, "two specials");
    constexpr int SW = NW >= 2 ? NW - 2 : 0; // the warp that finishes the special rows (not the manager warp)
    const bool sp1 = lane >= Q;               // lanes 0..Q-1: special 0, lanes Q..2Q-1: special 1
    const float sp_u = sp1 ? m.u[1] : m.u[0], sp_s0 = sp1 ? m.ss[1][0] : m.ss[0][0];
    const float sp_s1 = sp1 ? m.ss[1][1] : m.ss[0][1], sp_start = sp1 ? m.start_s[1] : m.start_s[0];
    const float* __restrict__ start_p = m.start_p + tid * R;

    float v[Q][R];
#pragma unroll
    for (int q = 0; q < Q; ++q)
#pragma unroll
        for (int r = 0; r < R; ++r) v[q][r] = pos_inf();

    // ---- sequence managers: identical protocol to the lockstep kernel (control blocks two steps ahead) ----
    unsigned long long mg_cur = 0, mg_nxt = 0;
    const unsigned long long* mg_ptr = sym64;
    uint32_t mg_pos = 0, mg_t = 0, mg_len = 0;
    auto next_symbol = [&]() -> uint32_t {
        const uint32_t sym = static_cast<uint32_t>(mg_cur >> ((mg_pos & 7u) * 8u)) & 0xffu;
        if ((++mg_pos & 7u) == 0) {
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
                col0 = next_symbol();
                started = true;
                one = mg_len == 1;
                if (mg_len > 1) {
                    col1 = next_symbol();
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
        if (lane < 8 && (fresh || (lane < Q && ((mask >> lane) & 1u)))) {
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

    // flags and TMEM columns of the coming step, fetched before the barrier that ends the current one
    uint4 fl, cA, cB;  // cA, cB: the symbols of sequences 0-3, 4-7
    // A[o_q][positions 1,2,3,4,0], B[o_q][position 0] (+2 ignored columns with the x8 load).  Fetched in PAIRS of sequences:
    // tcgen05.wait::ld retires every fetch in flight, so the way to give a fetch more than one sequence's arithmetic of
    // cover is to wait once per two sequences -- the pair (q, q+1) in use, the pair (q+2, q+3) in flight.
    constexpr int NE = Q >= 4 ? 4 : 2;
    uint32_t eR[NE][8];
    // B[o_q][positions 1..4]: a ring of three over q, fetched TWO sequences ahead (a shared-memory load under this
    // kernel's load takes longer than one sequence's arithmetic: one ahead left 0.9 short-scoreboard stalls per issue)
    float4 bR[3];
    const uint32_t sB_own = static_cast<uint32_t>(__cvta_generic_to_shared(sB)) + static_cast<uint32_t>(tid) * 16u;
    auto fetch_a = [&](uint32_t sym, uint32_t (&e)[8]) {
        if constexpr (X8) tmem_ld8(tE + sym * CS, e);
        else tmem_ld6(tE + sym * CS, e);
    };
    auto fetch_b = [&](uint32_t sym, float4& b) { b = lds128<0>(sB_own + sym * static_cast<uint32_t>(TPG * 16)); };
    auto prefetch_step = [&](const LsCtl& nc) {
        fl = *reinterpret_cast<const uint4*>(&nc.init_mask);
        cA = *reinterpret_cast<const uint4*>(&nc.col[0]);
        cB = *reinterpret_cast<const uint4*>(&nc.col[4]);
        fetch_a(cA.x, eR[0]);
        if constexpr (Q > 1) fetch_a(cA.y, eR[1]);
        fetch_b(cA.x, bR[0]);
        if constexpr (Q > 1) fetch_b(cA.y, bR[1]);
    };
    prefetch_step(ctl[0]);

    uint32_t tau = 0;
    for (;;) {
        const uint32_t init_mask = fl.x, fin_mask = fl.y;
        if (fl.z == 0) break; // no sequence takes part in this step: the CTA is done
        const uint32_t bo = tau & 1u, bi = bo ^ 1u; // buffers written / read in this step
        const uint32_t cols[8] = {cA.x, cA.y, cA.z, cA.w, cB.x, cB.y, cB.z, cB.w};

        // ---- managers: control block of step tau + 2 ----
        if (warp == MW) {
            LsCtl& nx = ctl[(tau + 2u) & 3u];
            bool cont = false, last = false;
            uint32_t colv = 0;
            if (lane < Q && mg_t + 1u < mg_len) {
                cont = true;
                colv = next_symbol();
                last = mg_t + 2u == mg_len;
                ++mg_t;
            }
            const uint32_t contm = __ballot_sync(0xffffffffu, cont);
            const uint32_t lastm = __ballot_sync(0xffffffffu, last);
            if (lane < 8) nx.col[lane] = colv;
            if (lane == 0) {
                nx.init_mask = 0;
                nx.fin_mask = lastm;
                nx.any_act = contm;
            }
        }

        // ---- special rows of step tau from the partial minima and special scores of step tau - 1: ONE warp, lane
        //      (s, q) finishes special row s of sequence q (the minimum over the NW warp partials is a short serial
        //      loop per lane instead of a REDUX per sequence on eight different warps) ----
        if (warp == SW && lane < SV_CHAIN_NSP * Q) {
            const int q = lane & (Q - 1);
            float mm = s_red[bi][0][q];
#pragma unroll
            for (int w = 1; w + 1 < NW; w += 2) mm = fminf(mm, fminf(s_red[bi][w][q], s_red[bi][w + 1][q]));
            if constexpr ((NW & 1) == 0) mm = fminf(mm, s_red[bi][NW - 1][q]);
            const float e = s_he[(sp1 ? kChainHeStride : 0) + ctl[tau & 3u].col[q]];
            float x = relax1(e, sp_u, mm);
            x = fminf(x, fminf(relax1(e, sp_s0, s_vs[bi][0][q]), relax1(e, sp_s1, s_vs[bi][1][q])));
            if ((init_mask >> q) & 1u) x = e + sp_start;
            s_vs[bo][sp1 ? 1 : 0][q] = x;
        }

        // ---- chain positions, one group of QV sequences at a time.  Idle sequence slots are relaxed like the
        //      others (their scores are never read: cheaper than a branch per sequence); everything a sequence
        //      needs from other threads arrives as ONE vector load per group: the hub score and the band
        //      predecessor of this thread's first position ----
        auto seq_step = [&](auto qc, float hub, float pv, float& red) {
            constexpr int q = decltype(qc)::value;
            uint32_t(&ec)[8] = eR[q % NE];
            const float4 bc = bR[q % 3];
            if constexpr ((q & 1) == 0) {
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                asm volatile("" : "+r"(ec[0]), "+r"(ec[1]), "+r"(ec[2]), "+r"(ec[3]), "+r"(ec[4]), "+r"(ec[5]));
                if constexpr (q + 1 < Q) {
                    uint32_t(&e1)[8] = eR[(q + 1) % NE];
                    asm volatile("" : "+r"(e1[0]), "+r"(e1[1]), "+r"(e1[2]), "+r"(e1[3]), "+r"(e1[4]), "+r"(e1[5]));
                }
                if constexpr (q + 2 < Q) fetch_a(cols[q + 2 < Q ? q + 2 : 0], eR[(q + 2) % NE]);
                if constexpr (q + 3 < Q) fetch_a(cols[q + 3 < Q ? q + 3 : 0], eR[(q + 3) % NE]);
            }
            if constexpr (q + 2 < Q) fetch_b(cols[q + 2 < Q ? q + 2 : 0], bR[(q + 2) % 3]);
            {
                // ec = A = fl(E + w_pred) of positions 1,2,3,4,0 and B = fl(E + w_hub) of position 0; bc = B of positions
                // 1..4.  v'[p] = min(fl(A[p] + v[p-1]), fl(B[p] + v[hub])): the oracle's fl(fl(E + T') + v), inner add tabulated.
                // Band predecessors of (1,2) are (0,1), of (3,4) are (2,3).
                const f32x2 A12 = pack2(__uint_as_float(ec[0]), __uint_as_float(ec[1]));
                const f32x2 A34 = pack2(__uint_as_float(ec[2]), __uint_as_float(ec[3]));
                const f32x2 hh = pack2(hub, hub);
                const f32x2 a12 = add2(A12, pack2(v[q][0], v[q][1]));
                const f32x2 a34 = add2(A34, pack2(v[q][2], v[q][3]));
                const f32x2 b12 = add2(pack2(bc.x, bc.y), hh);
                const f32x2 b34 = add2(pack2(bc.z, bc.w), hh);
                const float a0 = __uint_as_float(ec[4]) + pv, b0 = __uint_as_float(ec[5]) + hub;
                float a1, a2, a3, a4, b1, b2, b3, b4;
                unpack2(a12, a1, a2);
                unpack2(a34, a3, a4);
                unpack2(b12, b1, b2);
                unpack2(b34, b3, b4);
                v[q][0] = fminf(a0, b0);
                v[q][1] = fminf(a1, b1);
                v[q][2] = fminf(a2, b2);
                v[q][3] = fminf(a3, b3);
                v[q][4] = fminf(a4, b4);
            }
            float mq = fminf(v[q][0], fminf(v[q][1], v[q][2]));
            mq = fminf(mq, fminf(v[q][3], v[q][4]));
            red = warp_min_redux<NN>(mq);
        };
        auto group_step = [&](auto gc) {
            constexpr int g = decltype(gc)::value;
            float hub[QV], pv[QV], red[QV], top[QV];
            lds_vec<QV>(&s_vs[bi][0][g * QV], hub);
            lds_vec<QV>(&s_top[bi][g][tid * QV], pv);
            seq_step(std::integral_constant<int, g * QV>{}, hub[0], pv[0], red[0]);
            if constexpr (QV > 1) seq_step(std::integral_constant<int, g * QV + (QV > 1 ? 1 : 0)>{}, hub[QV > 1 ? 1 : 0], pv[QV > 1 ? 1 : 0], red[QV > 1 ? 1 : 0]);
            if constexpr (QV > 2) {
                seq_step(std::integral_constant<int, g * QV + (QV > 2 ? 2 : 0)>{}, hub[QV > 2 ? 2 : 0], pv[QV > 2 ? 2 : 0], red[QV > 2 ? 2 : 0]);
                seq_step(std::integral_constant<int, g * QV + (QV > 2 ? 3 : 0)>{}, hub[QV > 2 ? 3 : 0], pv[QV > 2 ? 3 : 0], red[QV > 2 ? 3 : 0]);
            }
#pragma unroll
            for (int j = 0; j < QV; ++j) top[j] = v[g * QV + j][R - 1];
            sts_vec<QV>(&s_top[bo][g][(tid + 1) * QV], top);
            if (lane == 0) sts_vec<QV>(&s_red[bo][warp][g * QV], red);
        };
        group_step(std::integral_constant<int, 0>{});
        if constexpr (NG > 1) {
            // an idle second group (the tail of a batch, see refill) is skipped; the emission fetch issued for its first
            // sequence is retired so that the next step's fetch into the same registers cannot be overtaken
            if ((fl.z & 0xf0u) != 0) {
                group_step(std::integral_constant<int, (NG > 1 ? 1 : 0)>{});
            } else {
                // sequence 2 issued the fetches for sequences 4 and 5: wait with their registers still allocated (an
                // asynchronous tcgen05.ld must never land in registers the compiler has handed to something else)
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                asm volatile("" ::"r"(eR[0][0]), "r"(eR[0][1]), "r"(eR[0][2]), "r"(eR[0][3]), "r"(eR[0][4]), "r"(eR[0][5]), "r"(eR[0][6]), "r"(eR[0][7]));
                asm volatile("" ::"r"(eR[1][0]), "r"(eR[1][1]), "r"(eR[1][2]), "r"(eR[1][3]), "r"(eR[1][4]), "r"(eR[1][5]), "r"(eR[1][6]), "r"(eR[1][7]));
            }
        }
        // ---- rare: first step of a sequence, v_0 = E[o_0] + start replaces the relaxed scores (a real
        //      branch around everything, so the hot path carries no predicated loads) ----
        if (init_mask != 0) {
            auto init_fix = [&](auto qc) {
                constexpr int q = decltype(qc)::value;
                if (!((init_mask >> q) & 1u)) return;
                const float* __restrict__ em = m.emis_p + static_cast<size_t>(cols[q]) * n_pos + tid * R;
#pragma unroll
                for (int r = 0; r < R; ++r) v[q][r] = em[r] + start_p[r];
                float mq = fminf(v[q][0], fminf(v[q][1], v[q][2]));
                mq = fminf(mq, fminf(v[q][3], v[q][4]));
                mq = warp_min_redux<NN>(mq);
                s_top[bo][q / QV][(tid + 1) * QV + q % QV] = v[q][R - 1];
                if (lane == 0) s_red[bo][warp][q] = mq;
            };
            init_fix(std::integral_constant<int, 0>{});
            if constexpr (Q > 1) init_fix(std::integral_constant<int, 1>{});
            if constexpr (Q > 2) {
                init_fix(std::integral_constant<int, 2>{});
                init_fix(std::integral_constant<int, 3>{});
            }
            if constexpr (Q > 4) {
                init_fix(std::integral_constant<int, 4>{});
                init_fix(std::integral_constant<int, 5>{});
                init_fix(std::integral_constant<int, 6>{});
                init_fix(std::integral_constant<int, 7>{});
            }
        }

        // control block of step tau + 1 (published during step tau - 1) and its first emission scores
        prefetch_step(ctl[(tau + 1u) & 3u]);
        __syncthreads();

        // ---- rare path: write out finished sequences, hand out new ones ----
        if (fin_mask != 0) {
            auto write_out = [&](auto qc) {
                constexpr int q = decltype(qc)::value;
                if (!((fin_mask >> q) & 1u)) return;
                const uint32_t sid = mgr[q].sid;
                float bv = pos_inf();
                int bidx = 0x7fffffff;
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const int o = m.orig_of_pos[tid * R + r];
                    if (o < 0) continue;
                    const float x = v[q][r];
                    if (out != nullptr) out[static_cast<size_t>(sid) * m.n + o] = x;
                    if (x < bv || (x == bv && o < bidx)) {
                        bv = x;
                        bidx = o;
                    }
                }
                if (tid < m.nsp) {
                    const float x = s_vs[bo][tid][q];
                    const int o = tid == 1 ? m.special[1] : m.special[0];
                    if (out != nullptr) out[static_cast<size_t>(sid) * m.n + o] = x;
                    if (x < bv || (x == bv && o < bidx)) {
                        bv = x;
                        bidx = o;
                    }
                }
                if (best != nullptr || arg != nullptr) {
#pragma unroll
                    for (int d = 16; d > 0; d >>= 1) {
                        const float ov = __shfl_xor_sync(0xffffffffu, bv, d);
                        const int oi = __shfl_xor_sync(0xffffffffu, bidx, d);
                        if (ov < bv || (ov == bv && oi < bidx)) {
                            bv = ov;
                            bidx = oi;
                        }
                    }
                    if (lane == 0) {
                        s_bv[warp] = bv;
                        s_bi[warp] = bidx;
                    }
                    __syncthreads();
                    if (warp == 0) {
                        bv = lane < NW ? s_bv[lane] : pos_inf();
                        bidx = lane < NW ? s_bi[lane] : 0x7fffffff;
#pragma unroll
                        for (int d = 16; d > 0; d >>= 1) {
                            const float ov = __shfl_xor_sync(0xffffffffu, bv, d);
                            const int oi = __shfl_xor_sync(0xffffffffu, bidx, d);
                            if (ov < bv || (ov == bv && oi < bidx)) {
                                bv = ov;
                                bidx = oi;
                            }
                        }
                        if (lane == 0) {
                            if (best != nullptr) best[sid] = bv;
                            // every score +inf: no best state (same convention as the other kernels)
                            if (arg != nullptr) arg[sid] = bv == pos_inf() ? -1 : bidx;
                        }
                    }
                    __syncthreads();
                }
            };
            write_out(std::integral_constant<int, 0>{});
            if constexpr (Q > 1) write_out(std::integral_constant<int, 1>{});
            if constexpr (Q > 2) {
                write_out(std::integral_constant<int, 2>{});
                write_out(std::integral_constant<int, 3>{});
            }
            if constexpr (Q > 4) {
                write_out(std::integral_constant<int, 4>{});
                write_out(std::integral_constant<int, 5>{});
                write_out(std::integral_constant<int, 6>{});
                write_out(std::integral_constant<int, 7>{});
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

// Instantiations: packed adds only (the scalar-add form, PK = false, measured 2.4 % slower on the 8-sequence kernel and
// is kept in the template as the readable statement of the same arithmetic).
template <int T, int QQ> const void* chain_fn_tq(bool nn) {
    return nn ? (const void*)sv_chain_kernel<T, QQ, true, true> : (const void*)sv_chain_kernel<T, QQ, false, true>;
}
template <int T> const void* chain_fn_t(int q, bool nn) {
    switch (q) {
    case 8: return chain_fn_tq<T, 8>(nn);
    case 4: return chain_fn_tq<T, 4>(nn);
    case 2: return chain_fn_tq<T, 2>(nn);
    case 1: return chain_fn_tq<T, 1>(nn);
    default: return nullptr;
    }
}
const void* chain_fn(int threads, int q, bool nn) {
    if (threads == 512) return chain_fn_t<512>(q, nn);
    if (threads == 256) return chain_fn_t<256>(q, nn);
    if (threads == 128) return chain_fn_t<128>(q, nn);
    return nullptr;
}

} // namespace

// Tensor memory is allocated once per CTA for its lifetime: at most 512 / tmem_cols CTAs may share an
// SM, otherwise a CTA would spin in tcgen05.alloc until a neighbour exits.  The dynamic shared-memory
// request is padded so that the hardware never co-schedules more than that.
int sv_chain_prepare(int threads, int q, bool nonneg, int emits, size_t* smem_out, int* ctas_per_sm_out) {
    const void* fn = chain_fn(threads, q, nonneg);
    if (fn == nullptr) return cudaErrorInvalidValue;
    const int want = threads == 512 ? 1 : threads == 256 ? 2 : 4;
    cudaFuncAttributes fa;
    cudaError_t e = cudaFuncGetAttributes(&fa, fn);
    if (e != cudaSuccess) return e;
    int dev = 0, smem_sm = 0, smem_optin = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
    cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    // dynamic shared memory = the fl(E + w_hub) table of positions 1..4 (one float4 per symbol and thread); then grow the
    // request until the occupancy calculator agrees
    size_t dyn = static_cast<size_t>(emits) * threads * 16;
    if (fa.sharedSizeBytes + dyn > static_cast<size_t>(smem_optin)) return cudaErrorLaunchOutOfResources;
    int occ = 0;
    for (;;) {
        e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(dyn));
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, threads, dyn);
        if (e != cudaSuccess) return e;
        if (occ <= want) break;
        const size_t next = dyn + 4096;
        if (fa.sharedSizeBytes + next > static_cast<size_t>(smem_optin)) return cudaErrorLaunchOutOfResources;
        dyn = next;
    }
    (void)smem_sm;
    if (occ < 1) return cudaErrorLaunchOutOfResources;
    *smem_out = dyn;
    *ctas_per_sm_out = occ;
    return cudaSuccess;
}

int sv_chain_launch(int threads, int q, const SvChainDev& m, const SvSeqsDev& s, float* d_out, float* d_best,
                    int32_t* d_arg, unsigned* d_counter, int grid, size_t smem, cudaStream_t st) {
    const void* fn = chain_fn(threads, q, m.nonneg != 0);
    if (fn == nullptr || s.sym_bytes != 1) return cudaErrorInvalidValue;
    void* args[] = {(void*)&m, (void*)&s, (void*)&d_out, (void*)&d_best, (void*)&d_arg, (void*)&d_counter};
    cudaError_t e = cudaLaunchKernel(fn, dim3(grid), dim3(threads), args, smem, st);
    return e != cudaSuccess ? e : cudaGetLastError();
}
