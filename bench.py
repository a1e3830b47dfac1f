#!/usr/bin/env python
"""Benchmark of the (min,+) Viterbi hot path on B200 -- contract in DESIGN.md section "Measurement".

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]

One "step" = one pass of the plain (min,+) Viterbi recursion over one batch of sequences:
workload `synthetic2405` (BASELINE.json configs[4]) = a seeded chmm_gen-style 2405-state HMM
(nnz 7215, 20 symbols) x 4096 uniform random sequences of 7000 symbols PER GPU (weak scaling,
sequences are independent; the HMM is replicated).  Metric = state-updates/s = S*T*N / seconds.

  value     kernel throughput with the model and the packed symbols resident in HBM
  e2e       same metric through the host-buffer C-ABI call sv_viterbi_batch_u8 (pinned host
            symbols -> device, kernel, full final vectors + best scores -> pinned host)
  roofline  dominant kernel (the persistent batched step) against the shared-memory crossbar
            (128 B/clk/SM), the lower of its two ceilings; the non-tensor FP32 issue ceiling is
            reported beside it (`fp32_issue`).  (min,+) is not an MMA contraction and the working
            set lives in shared memory / registers, so neither HBM nor the tensor pipe bounds it
  cpu_baseline  the oracle port (oracle/sv_oracle.c, OpenMP over sequences) on the host cores

`--impl reference` times the reference's own CPU implementation (oracle/_ref = its
GraphBLAS_impl sources compiled unmodified on the GraphBLAS shim; the oracle port if that was
not built) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "state_updates_per_s"
UNIT = "state-updates/s"

WORKLOADS = {
    # name: (model, nseq per GPU, seq len)
    "synthetic2405": dict(model="synthetic", states=2405, nseq=4096, T=7000),
    "profile2405": dict(model="2405.chmmb", nseq=4096, T=7000),
    "small": dict(model="synthetic", states=300, nseq=256, T=500),
}


def load_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return json.loads(p.read_text()), "measured"
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


def ncu_traffic(kernel_substr: str, workload: str):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` summary
    (profiles/*_ncu_summary.txt, written by tools/ncu_summary.py); None if there is none for this workload."""
    if workload != "synthetic2405":
        return None, None
    best = None
    for f in sorted((ROOT / "profiles").glob("*_ls_*ncu_summary.txt")):
        if "profile2405" in f.name:
            continue
        rd = wr = None
        ok = False
        for ln in f.read_text().splitlines():
            parts = ln.split("\t")
            if parts[0] == "Kernel Name":
                ok = kernel_substr in parts[-1]
            if len(parts) == 3 and ok:
                scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(parts[1])
                if parts[0] == "dram__bytes_read.sum" and scale:
                    rd = float(parts[2]) * scale
                if parts[0] == "dram__bytes_write.sum" and scale:
                    wr = float(parts[2]) * scale
        if rd is not None and wr is not None:
            best = (rd + wr, f.name)
    return best if best else (None, None)


def make_workload(name: str, rank: int):
    from spec_viterbi_b200 import api, generators
    w = WORKLOADS[name]
    if w["model"] == "synthetic":
        hmm = generators.random_hmm(states_num=w["states"], seed=w["states"])
    else:
        hmm = api.read_HMM(str(ROOT / "tests" / "golden" / "models" / w["model"]))
    syms, offs = generators.random_sequences(w["nseq"], w["T"], emit_range=int(hmm.emit_num),
                                             seed=7000 + rank, dtype=np.uint8)
    return hmm, syms, offs


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.02)

    def __enter__(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def physical_index(local_rank: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


def cpu_port_baseline(hmm, syms, offs, seconds_target=12.0):
    """Oracle port on all host cores over a bounded sample of the same workload."""
    from oracle import pyoracle
    oracle = pyoracle.Oracle()
    cores = os.cpu_count() or 1
    n, T = int(hmm.states_num), int(offs[1] - offs[0])
    per_seq = n * T / 1.0e8  # ~1e8 state-updates/s/thread (measured in the build container)
    nseq = int(max(cores, min(offs.size - 1, cores * max(1, round(seconds_target / per_seq)))))
    s64 = syms[: int(offs[nseq])].astype(np.uint64)
    o = offs[: nseq + 1]
    oracle.plain_batch(hmm, s64[: int(o[min(nseq, cores)])], o[: min(nseq, cores) + 1], threads=cores)  # warm
    t0 = time.perf_counter()
    oracle.plain_batch(hmm, s64, o, threads=cores)
    dt = time.perf_counter() - t0
    return {"value": nseq * T * n / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"first {nseq} of the rank-0 sequences ({T} symbols each), oracle/sv_oracle.c, "
                      f"OpenMP over sequences, {dt:.2f} s"}


def run_reference(args):
    """Reference arm: the reference's CPU GraphBLAS_impl (one run_Viterbi call per sequence,
    bench_Viterbi.h:51-57) on all host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pyoracle
    hmm, syms, offs = make_workload(args.workload, 0)
    cores = os.cpu_count() or 1
    n, T = int(hmm.states_num), int(offs[1] - offs[0])
    use_ref = pyoracle.REF_SO.exists()
    if use_ref:
        impl, kind, rate = pyoracle.Reference(), "reference", 1.0e7
    else:
        impl, kind, rate = pyoracle.Oracle(), "port", 1.0e8
    # ~6 s of CPU work per step
    nseq = int(max(cores, min(offs.size - 1, cores * max(1, round(6.0 * rate / (n * T))))))
    s64 = syms[: int(offs[nseq])].astype(np.uint64)
    o = offs[: nseq + 1]
    for _ in range(max(0, args.warmup)):
        impl.plain_batch(hmm, s64[: int(o[cores])], o[: cores + 1], threads=cores)
    times = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        impl.plain_batch(hmm, s64, o, threads=cores)
        times.append(time.perf_counter() - t0)
    dt = float(np.mean(times))
    value = nseq * T * n / dt
    sample = (f"{nseq} sequences x {T} symbols per step on {cores} host threads, "
              + ("reference GraphBLAS_impl.cpp on oracle/grb_shim (SuiteSparse absent)" if use_ref
                 else "oracle port (oracle/_ref not built)"))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, hmm, offs, 1),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_line(line)


def workload_config(name, hmm, offs, n_gpus):
    nseq = offs.size - 1
    T = int(offs[1] - offs[0])
    return {"workload": name, "states": int(hmm.states_num), "transitions": int(hmm.trans_num),
            "symbols": int(hmm.emit_num), "sequences_per_gpu": int(nseq), "sequence_length": T,
            "path": "plain (min,+) step, batched sequences", "parallelism": f"sequence-sharded x{n_gpus}",
            "l2": "flushed between timed steps (256 MiB write)"}


def run_b200(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this backend has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from spec_viterbi_b200 import api, _lib
    hmm, syms, offs = make_workload(args.workload, rank)
    n, nseq, T = int(hmm.states_num), offs.size - 1, int(offs[1] - offs[0])
    model = api.Model(hmm, device=local_rank)
    if args.kernel == "generic":
        model.set_kernel(_lib.KERNEL_GENERIC)
    elif args.kernel == "es":
        model.set_kernel(_lib.KERNEL_ES)
    elif args.kernel == "ls":
        model.set_kernel(_lib.KERNEL_LS)
    info = model.info()
    seqs = model.upload(syms, offs)
    d_out = torch.empty((nseq, n), dtype=torch.float32, device=dev)
    d_best = torch.empty(nseq, dtype=torch.float32, device=dev)
    d_arg = torch.empty(nseq, dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()

    def step():
        model.viterbi_device(seqs, d_out.data_ptr(), d_best.data_ptr(), d_arg.data_ptr(), stream.cuda_stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step()
    barrier()

    # ---- timed region: K steps, CUDA events around each step on the launching stream ----
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(physical_index(local_rank)) as clocks:
        barrier()
        for a, b in evs:
            flush.fill_(1)  # L2 flush, outside the event pair
            a.record(stream)
            step()
            b.record(stream)
        barrier()
    step_ms = [a.elapsed_time(b) for a, b in evs]
    total_ms = float(sum(step_ms))
    launches = model.last_launches() * args.steps

    # ---- e2e: host buffers through the C ABI (pinned symbols in, vectors + best out) ----
    h_syms = torch.from_numpy(syms).pin_memory()
    h_out = torch.empty((nseq, n), dtype=torch.float32).pin_memory()
    h_best = torch.empty(nseq, dtype=torch.float32).pin_memory()
    h_arg = torch.empty(nseq, dtype=torch.int32).pin_memory()
    np_syms, np_out, np_best, np_arg = h_syms.numpy(), h_out.numpy(), h_best.numpy(), h_arg.numpy()

    def e2e_step():
        model.viterbi_batch(np_syms, offs, out=np_out, best=np_best, arg=np_arg)

    e2e_steps = max(2, min(args.steps, 5))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()  # synchronous: returns when the results are in host memory
    barrier()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    h2d = int(syms.nbytes + offs.nbytes + 4 * nseq)
    d2h = int(np_out.nbytes + np_best.nbytes + np_arg.nbytes)

    # parity spot check of what was timed (not part of the timing): rank 0, a few sequences
    if rank == 0:
        from oracle import pyoracle
        oracle = pyoracle.Oracle()
        k = 3
        exp = oracle.plain_batch(hmm, syms[: int(offs[k])].astype(np.uint64), offs[: k + 1], threads=k)
        got = d_out[:k].cpu().numpy()
        if not (np.array_equal(exp, got) and np.array_equal(exp, np_out[:k])):
            raise SystemExit("bench.py: GPU results differ from the oracle -- number is invalid")

    # ---- second headline figure: spec precompute ms (level 2: emits^2 pair handlers), tuple-sharded
    #      over the ranks into one caller-owned table, completed by an all-gather of the shards ----
    from spec_viterbi_b200 import sharding
    E = int(hmm.emit_num)
    ntup = E * E
    table = torch.empty((ntup, n, n), dtype=torch.float32, device=dev)
    def timed_build(tb, te, gather):
        best_b, best_g = 1e30, 1e30
        for _ in range(3):
            barrier()
            t0 = time.perf_counter()
            sp = api.Spec(model, 2, _lib.SPEC_AUTO, tb, te, d_table=table[tb:te].data_ptr())  # synchronous
            t1 = time.perf_counter()
            if gather:
                sharding.all_gather_table(table, ntup)
            barrier()
            t2 = time.perf_counter()
            sp.close()
            best_b, best_g = min(best_b, t1 - t0), min(best_g, t2 - t1)
        return best_b, best_g
    # (a) every GPU builds the whole table locally: no traffic; (b) N > 1: tuple shards + all-gather
    rep_build_s, _ = timed_build(0, ntup, False)
    tb, te = sharding.shard_tuples(ntup, world, rank)
    spec_build_s, spec_gather_s = timed_build(tb, te, True) if world > 1 else (rep_build_s, 0.0)
    del table

    # ---- max over ranks ----
    t = torch.tensor([total_ms, e2e_s, spec_build_s, spec_gather_s, rep_build_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # gather of final scores (north star: NCCL only to gather results)
        gathered = [torch.empty_like(d_best) for _ in range(world)]
        dist.all_gather(gathered, d_best)
    total_ms, e2e_s, spec_build_s, spec_gather_s, rep_build_s = (float(x) for x in t)
    updates_per_step = float(nseq) * T * n * world
    value = updates_per_step * args.steps / (total_ms * 1e-3)
    e2e_value = updates_per_step / e2e_s

    if rank == 0:
        peaks, peak_kind = load_peaks()
        clk = clocks.summary()
        sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
        mhz = clk["sm_mhz"] or peaks.get("sm_max_mhz", 1965.0)
        relax = float(nseq) * (T - 1) * int(info.nnz)          # relaxations per launch (per GPU)
        kernel_s = float(np.mean(step_ms)) * 1e-3
        # algorithmic ops: one add + one min per stored transition; peak = FP32 issue
        # bound of a (min,+) MAC = 1 FADD + 1/2 FMNMX3 = 1.5 slots on 128 lanes/SM at the SM clock
        # measured during the run
        achieved = 2.0 * relax / kernel_s / 1e12
        peak = 2.0 * sm_count * 128 * mhz * 1e6 / 1.5 / 1e12
        traffic, traffic_src = ncu_traffic("sv_ls_kernel", args.workload) if info.kernel == _lib.KERNEL_LS else (None, None)
        xbar_bytes = 4.0 * relax + 4.0 * nseq * T * n   # 4 B gathered per relaxation + 4 B stored per state update
        xbar_peak = sm_count * 128 * mhz * 1e6          # shared-memory crossbar: 128 B/clk/SM at the measured SM clock
        # K2 has two ceilings: FP32 issue and the shared-memory crossbar.  The crossbar is the
        # LOWER one (9.3e12 vs 2.5e13 relaxations/s) and the one ncu shows saturating, so it is the roofline
        # reported at the top level; the FP32-issue figure sits beside it.  Neither "hbm" nor "tensor"
        # bounds this kernel: its HBM traffic is the symbol stream in and the final vectors out.
        roofline = {"bound": "smem", "achieved": xbar_bytes / kernel_s / 1e9, "peak": xbar_peak / 1e9, "unit": "GB/s",
                    "frac": xbar_bytes / kernel_s / xbar_peak, "traffic": traffic, "traffic_source": traffic_src,
                    "kernel": {_lib.KERNEL_ES: "sv_es_kernel", _lib.KERNEL_LS: "sv_ls_kernel"}.get(info.kernel, "sv_plain_generic_kernel"),
                    "relaxations_per_launch": relax, "relax_per_clk_per_sm": relax / kernel_s / (sm_count * mhz * 1e6),
                    "peak_source": f"shared-memory crossbar, {sm_count} SMs x 128 B/clk x {mhz:.0f} MHz (NVML median during the "
                                   f"run); algorithmic bytes = 4 B gathered per relaxation + 4 B stored per state update; "
                                   f"MEASURED_PEAKS.json ({peak_kind}) holds only HBM/bf16 peaks, neither of which bounds this kernel",
                    "fp32_issue": {"achieved_TFLOPs": achieved, "peak_TFLOPs": peak, "frac": achieved / peak,
                                   "peak_source": "128 lanes/SM at the same clock / 1.5 issue slots per (min,+) MAC "
                                                  "(1 FADD + 1/2 FMNMX3); algorithmic ops = 2 per relaxation"},
                    "hbm_algorithmic_gbs": (syms.nbytes + d_out.numel() * 4) / kernel_s / 1e9,
                    "hbm_peak_gbs": peaks.get("hbm_gbs")}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args.workload, hmm, offs, world),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "sv_viterbi_batch_u8 (host u8 symbols -> final vectors + best score/state)"},
            "gpu_launches": int(launches), "clocks": clk, "roofline": roofline,
            "kernel": {"name": roofline["kernel"], "es_threads": int(info.es_threads),
                       "es_seqs_per_cta": int(info.es_seqs_per_cta), "es_smem_bytes": int(info.es_smem_bytes),
                       "ls_seqs_per_cta": int(info.ls_seqs_per_cta), "ls_smem_bytes": int(info.ls_smem_bytes),
                       "ls_tmem_cols": int(info.ls_tmem_cols),
                       "ls_gather_wavefronts": int(info.ls_gather_wavefronts),
                       "ls_gather_ideal": int(info.ls_gather_ideal),
                       "step_ms": step_ms},
        }
        table_gb = ntup * n * n * 4 / 1e9
        sharded_ms = (spec_build_s + spec_gather_s) * 1e3
        line["spec_precompute"] = {
            "level": 2, "handlers": ntup, "table_GB": table_gb,
            "precompute_ms": min(rep_build_s * 1e3, sharded_ms) if world > 1 else rep_build_s * 1e3,
            "strategy": "replicated" if world == 1 or rep_build_s * 1e3 <= sharded_ms else "tuple-sharded + all-gather",
            "replicated_build_ms": rep_build_s * 1e3,
            "sharded_build_ms": spec_build_s * 1e3 if world > 1 else None,
            "sharded_allgather_ms": spec_gather_s * 1e3 if world > 1 else None,
            "kernel": "pair_rows_kernel (K3'')",
            "write_GBs_per_gpu": table_gb / rep_build_s,
            "frac_of_hbm": table_gb / rep_build_s / peaks.get("hbm_gbs", 6546.2),
            "timing": "wall clock around the synchronous sv_spec_create into a caller-owned table (every GPU holds "
                      "the whole table afterwards), min of 3, max over ranks"}
        if not args.no_cpu:
            line["cpu_baseline"] = cpu_port_baseline(hmm, syms, offs)
        emit_line(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_JSON_OUT = None


def emit_line(line) -> None:
    out = _JSON_OUT if _JSON_OUT is not None else sys.stdout
    print(json.dumps(line), file=out, flush=True)


def claim_stdout() -> None:
    """stdout carries ONE JSON line: everything else written to fd 1 (NCCL prints its version banner there from C)
    is sent to stderr; the JSON line goes to a private duplicate of the original stdout."""
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="synthetic2405", choices=sorted(WORKLOADS))
    ap.add_argument("--kernel", default="auto", choices=["auto", "generic", "es", "ls"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
