/* sv_probe_b200 -- micro-benchmarks behind the roofline figures of spec_viterbi_b200.
 *
 * A SEPARATE library (spec_viterbi_b200/libsvprobe_b200.so, built by spec_viterbi_b200/build.py from
 * csrc/sv_probe.cu): measurement infrastructure, not part of the drop-in boundary.  The product library
 * neither links nor loads it.  Every function returns 0 on success (1 = bad argument, 2 = CUDA error;
 * message in sv_probe_last_error()).
 */
#ifndef SV_PROBE_B200_H
#define SV_PROBE_B200_H

#ifdef __cplusplus
extern "C" {
#endif

const char* sv_probe_last_error(void);

/* FP32 pipe probe used to fix the non-tensor roofline: runs `iters` dependent-chain iterations of the
 * chosen instruction mix on every SM and returns lane-instructions/s.
 * mix: 0 = FADD, 1 = FMNMX (2-input min), 2 = FMNMX3 (3-input min), 3 = 2 FADD + 1 FMNMX3,
 *      4 = 4 FADD + 1 FMNMX3 (the oracle-order relaxation pair), 5 = LDS.32 conflict-free + 2 FADD + FMNMX,
 *      6 = FADD2 (packed add.rn.f32x2, counted as two adds), 7 = 2 FADD2 + 1 FMNMX3 (the same pair, packed) */
int sv_probe_fp32(int device, int mix, int iters, double* lane_instr_per_s, double* sm_clock_mhz);

/* Tensor-memory probe: fills all 512 TMEM columns of every SM through tcgen05.st, reads them back with
 * run-time (warp-uniform) column addresses -- the access pattern of the emission lookup in the step
 * kernels -- and reports mismatches and the tcgen05.ld read rate. */
int sv_probe_tmem(int device, int iters, unsigned* errors, double* bytes_per_clk_per_sm);

/* Shared-memory pipe probe under the lockstep kernel's access pattern (16 warps per SM, 8 loads in flight
 * per thread, per-lane register addresses that are bank-conflict free by construction).
 * mode: 1 = 3 LDS.128 : 1 STS.128 (the kernel's ratio), 2 = mode 1 + the relaxations (the whole inner loop).
 * Returns bytes (and lane loads) per clock and SM of the slowest SM.  Both modes were checked against ncu's
 * shared-memory wavefront counters (profiles/r02_smem_probe_ncu.csv). */
int sv_probe_smem(int device, int mode, int iters, double* bytes_per_clk_per_sm, double* loads_per_clk_per_sm);

#ifdef __cplusplus
}
#endif
#endif
