"""Instrumented-build helper (SV_NVCC_EXTRA=-DSV_LS_DEBUG): per-warp split-barrier wait cycles of the LS kernel."""
import sys, ctypes as C, numpy as np
sys.path.insert(0, '.')
from spec_viterbi_b200 import api, generators, _lib
hmm = generators.random_hmm(states_num=2405, seed=2405)
syms, offs = generators.random_sequences(4096, 2000, seed=7000, dtype=np.uint8)
m = api.Model(hmm)

out = m.viterbi_batch(syms, offs)
lib = _lib.lib()
buf = (C.c_ulonglong * (148 * 16 * 4))()
lib.sv_debug_ls_dump.argtypes = [C.POINTER(C.c_ulonglong)]
print('rc', lib.sv_debug_ls_dump(buf))
a = np.frombuffer(buf, dtype=np.uint64).reshape(148, 16, 4).astype(np.float64)
tot = a[:, :, 2].mean()
print('total clk', tot, 'steps', a[:, :, 3].mean())
print('waitA frac per warp', np.round(a[:, :, 0].mean(0) / tot, 3))
print('waitB frac per warp', np.round(a[:, :, 1].mean(0) / tot, 3))

tb = (C.c_ulonglong * (16 * 8 * 8))()
lib.sv_debug_ls_trace.argtypes = [C.POINTER(C.c_ulonglong)]
lib.sv_debug_ls_trace(tb)
t = np.frombuffer(tb, dtype=np.uint64).reshape(16, 8, 8).astype(np.int64)
t0 = t[:, 0, 0].min()
np.set_printoptions(linewidth=200)
for st in range(1, 4):
    print('step', st, 'events: top, -, preWaitB, postWaitB, arriveA, preWaitA, postWaitA, arriveB (clk rel.)')
    print((t[:, st, :] - t0))
