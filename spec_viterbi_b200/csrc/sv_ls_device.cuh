// Device helpers shared by the lockstep step kernels (sv_plain_ls.cu: one thread set per CTA;
// sv_plain_ls2.cu: two independent halves per CTA).  Internal to the library.
#pragma once
#include "sv_internal.h"

#include <type_traits>

namespace {

__device__ __forceinline__ float pos_inf() { return __int_as_float(0x7f800000); }

template <int... S> struct LShape {
    static constexpr int R = sizeof...(S);
    __host__ __device__ static constexpr int at(int r) {
        constexpr int v[] = {S...};
        return v[r];
    }
    __host__ __device__ static constexpr int base(int r) {
        int b = 0;
        for (int k = 0; k < r; ++k) b += at(k);
        return b;
    }
    __host__ __device__ static constexpr int total() { return base(R); }
};

// ---- tensor memory ----
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t ncols) {
    const uint32_t dst = static_cast<uint32_t>(__cvta_generic_to_shared(smem_slot));
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst), "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_st1(uint32_t taddr, float v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr), "r"(__float_as_uint(v))
                 : "memory");
}
// E[o][own rows of ranks 0..4] of one sequence: 5 consecutive TMEM columns at a run-time address.
__device__ __forceinline__ void tmem_ld5(uint32_t taddr, uint32_t (&r)[5]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(taddr)
                 : "memory");
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r[4]) : "r"(taddr + 4u) : "memory");
}
// v'[own row] of 4 sequences: 32-bit shared address + immediate (no generic-pointer arithmetic)
template <int IMM>
__device__ __forceinline__ void sts128(uint32_t addr, const float (&a)[4]) {
    asm volatile("st.shared.v4.f32 [%0+%1], {%2,%3,%4,%5};" ::"r"(addr), "n"(IMM), "f"(a[0]), "f"(a[1]), "f"(a[2]),
                 "f"(a[3])
                 : "memory");
}
// Per-step control block published by the manager lanes.
struct alignas(16) LsCtl {
    uint32_t col[8];    // TMEM column offset of each sequence's current symbol: symbol * RSTRIDE
    uint32_t init_mask; // sequences whose step is v_0 = E[o_0] + start
    uint32_t fin_mask;  // sequences for which this step is the last one
    uint32_t any_act;   // sequences that take part in this step (0 = CTA is done)
    uint32_t pad;
};

// Sequence-manager state of one sequence slot lives in REGISTERS of lane q of the manager warp:
// under load a shared-memory round trip takes hundreds of cycles (the crossbar queue is deep), and
// a manager that chases its state through shared memory falls a thousand cycles behind the other
// warps every step.  Only the sequence id (needed on the rare write-out path) is kept in memory.
struct LsMgr {
    uint32_t sid;
};

// fl(fl(E + T') + v) -- the two adds are never re-associated (GraphBLAS_impl.cpp:65,69)
__device__ __forceinline__ float relax1(float e, float w, float v) { return (e + w) + v; }

// Packed IEEE adds (sm_100+: add.rn.f32x2 -> FADD2): two independent round-to-nearest fp32 adds in ONE instruction
// on a 64-bit register pair.  Each half is an ordinary IEEE add -- fl(fl(E + T') + v) stays exactly the oracle's
// (GraphBLAS_impl.cpp:65,69) -- but a slot pair of 4 sequences issues 8 FADD2 instead of 16 FADD.
#ifndef SV_LS_PACKED
#define SV_LS_PACKED 1
#endif
typedef unsigned long long ls_f32x2;
__device__ __forceinline__ ls_f32x2 ls_pack2(float lo, float hi) {
    ls_f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void ls_unpack2(ls_f32x2 p, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p));
}
__device__ __forceinline__ ls_f32x2 ls_add2(ls_f32x2 a, ls_f32x2 b) {
    ls_f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

// One pair of slots for 4 sequences: acc = min(acc, x0, x1) is a single 3-input FMNMX3.
template <bool FIRST>
__device__ __forceinline__ void pair4(float w0, float w1, const float4 a, const float4 b, const float* e,
                                      float* acc) {
#if SV_LS_PACKED
    const ls_f32x2 e01 = ls_pack2(e[0], e[1]), e23 = ls_pack2(e[2], e[3]);
    const ls_f32x2 ww0 = ls_pack2(w0, w0), ww1 = ls_pack2(w1, w1);
    float p0, p1, p2, p3, q0, q1, q2, q3;
    ls_unpack2(ls_add2(ls_add2(e01, ww0), ls_pack2(a.x, a.y)), p0, p1);
    ls_unpack2(ls_add2(ls_add2(e23, ww0), ls_pack2(a.z, a.w)), p2, p3);
    ls_unpack2(ls_add2(ls_add2(e01, ww1), ls_pack2(b.x, b.y)), q0, q1);
    ls_unpack2(ls_add2(ls_add2(e23, ww1), ls_pack2(b.z, b.w)), q2, q3);
    const float x0 = fminf(p0, q0), x1 = fminf(p1, q1), x2 = fminf(p2, q2), x3 = fminf(p3, q3);
#else
    const float x0 = fminf(relax1(e[0], w0, a.x), relax1(e[0], w1, b.x));
    const float x1 = fminf(relax1(e[1], w0, a.y), relax1(e[1], w1, b.y));
    const float x2 = fminf(relax1(e[2], w0, a.z), relax1(e[2], w1, b.z));
    const float x3 = fminf(relax1(e[3], w0, a.w), relax1(e[3], w1, b.w));
#endif
    if (FIRST) {
        acc[0] = x0; acc[1] = x1; acc[2] = x2; acc[3] = x3;
    } else {
        acc[0] = fminf(acc[0], x0); acc[1] = fminf(acc[1], x1);
        acc[2] = fminf(acc[2], x2); acc[3] = fminf(acc[3], x3);
    }
}
template <bool FIRST>
__device__ __forceinline__ void single4(float w0, const float4 a, const float* e, float* acc) {
#if SV_LS_PACKED
    const ls_f32x2 ww0 = ls_pack2(w0, w0);
    float x0, x1, x2, x3;
    ls_unpack2(ls_add2(ls_add2(ls_pack2(e[0], e[1]), ww0), ls_pack2(a.x, a.y)), x0, x1);
    ls_unpack2(ls_add2(ls_add2(ls_pack2(e[2], e[3]), ww0), ls_pack2(a.z, a.w)), x2, x3);
#else
    const float x0 = relax1(e[0], w0, a.x), x1 = relax1(e[1], w0, a.y);
    const float x2 = relax1(e[2], w0, a.z), x3 = relax1(e[3], w0, a.w);
#endif
    if (FIRST) {
        acc[0] = x0; acc[1] = x1; acc[2] = x2; acc[3] = x3;
    } else {
        acc[0] = fminf(acc[0], x0); acc[1] = fminf(acc[1], x1);
        acc[2] = fminf(acc[2], x2); acc[3] = fminf(acc[3], x3);
    }
}

// Gather of one slot: volatile, so the software pipeline below is emitted in the order written
// (the relaxations float freely around the loads).  addr = absolute shared address of v[src] in
// buffer 0 + parity offset (a uniform register); the float4-group offset is an immediate.
template <int IMM>
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4+%5];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "r"(addr), "n"(IMM)
                 : "memory");
    return v;
}

#ifndef SV_LS_DEPTH
#define SV_LS_DEPTH 1 // slot pairs in flight beside the one being relaxed
#endif

// Gathered scores of one slot pair (4 sequences).
struct LsPair {
    float4 a, b;
};

// k / both are compile-time constants after unrolling (register indices stay static).  `eo(k)` returns
// the shared address (or offset) of slot k: a plain register array element, or one half of a register
// that packs two 16-bit offsets.
template <int GOFF, typename EO>
__device__ __forceinline__ void ls_gather(const EO& eo, uint32_t pofs, LsPair& d, int k, bool both) {
    d.a = lds128<GOFF>(eo(k) + pofs);
    if (both) d.b = lds128<GOFF>(eo(k + 1) + pofs);
}

// D slots (compile-time) of the register slice [BASE, BASE + D) for the 4 sequences of the float4
// group at byte offset GOFF, as straight-line code, slot PAIRS at a time with register double
// buffering: the gathers of pair p + 1 are issued before the relaxations of pair p, and before the
// last pair the FIRST pair of the next slice (NBASE >= 0: next rank, or rank 0 of the next group,
// at group offset NGOFF) is issued, so that only the first gather of a time step is exposed.
// `io` holds this slice's first pair on entry and the next slice's first pair on exit.
template <int GOFF, int BASE, int D, int NGOFF, int NBASE, bool NBOTH, int NSLOTS, typename EO>
__device__ __forceinline__ void ls_fixed(const float (&ew)[NSLOTS], const EO& eo, uint32_t pofs,
                                         uint32_t npofs, const float (&e)[4], float (&acc)[4], LsPair& io) {
    if constexpr (D == 0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[q] = pos_inf();
        if constexpr (NBASE >= 0) ls_gather<NGOFF>(eo, npofs, io, NBASE, NBOTH);
    } else {
        constexpr int NP = (D + 1) / 2;
        LsPair buf[NP + 1];
        buf[0] = io;
        // future pairs F = [pair 1, .., pair NP-1, first pair of the next slice]; before pair p is relaxed
        // F[0 .. p + DEPTH - 1] have been issued (DEPTH = SV_LS_DEPTH pairs in flight beside the one in use)
        auto issue = [&](int f) { // f = index into F, compile-time after unrolling
            if (f + 1 < NP)
                ls_gather<GOFF>(eo, pofs, buf[f + 1], BASE + 2 * f + 2, 2 * f + 3 < D);
            else if (f + 1 == NP && NBASE >= 0)
                ls_gather<NGOFF>(eo, npofs, buf[f + 1], NBASE, NBOTH);
        };
#pragma unroll
        for (int p = 0; p < NP; ++p) {
            if (p == 0) {
#pragma unroll
                for (int f = 0; f < SV_LS_DEPTH; ++f) issue(f);
            } else {
                issue(p + SV_LS_DEPTH - 1);
            }
            const int k = BASE + 2 * p;
            if (2 * p + 1 < D) {
                if (p == 0)
                    pair4<true>(ew[k], ew[k + 1], buf[p].a, buf[p].b, e, acc);
                else
                    pair4<false>(ew[k], ew[k + 1], buf[p].a, buf[p].b, e, acc);
            } else {
                if (p == 0)
                    single4<true>(ew[k], buf[p].a, e, acc);
                else
                    single4<false>(ew[k], buf[p].a, e, acc);
            }
        }
        if constexpr (NBASE >= 0) io = buf[NP];
    }
}

// Runs slots [0, d) of the register slice [BASE, BASE + SMAX).  d is warp-uniform and lies in
// [LO, HI]: a binary tree of uniform branches into straight-line code (a jump table costs a
// constant-bank load and an indirect branch per slice).
template <int GOFF, int BASE, int LO, int HI, int NGOFF, int NBASE, bool NBOTH, int NSLOTS, typename EO>
__device__ __forceinline__ void ls_run_slots(int d, const float (&ew)[NSLOTS], const EO& eo,
                                             uint32_t pofs, uint32_t npofs, const float (&e)[4], float (&acc)[4],
                                             LsPair& io) {
    static_assert(HI <= 8, "at most 8 slots per rank");
    if constexpr (LO == HI) {
        ls_fixed<GOFF, BASE, LO, NGOFF, NBASE, NBOTH, NSLOTS>(ew, eo, pofs, npofs, e, acc, io);
    } else {
        constexpr int MID = (LO + HI) / 2;
        if (d <= MID)
            ls_run_slots<GOFF, BASE, LO, MID, NGOFF, NBASE, NBOTH, NSLOTS>(d, ew, eo, pofs, npofs, e, acc, io);
        else
            ls_run_slots<GOFF, BASE, MID + 1, HI, NGOFF, NBASE, NBOTH, NSLOTS>(d, ew, eo, pofs, npofs, e, acc, io);
    }
}

} // namespace
