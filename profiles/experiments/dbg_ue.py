import sys, time, faulthandler
sys.path.insert(0, "."); sys.path.insert(0, "tests")
faulthandler.dump_traceback_later(25, exit=True)
import numpy as np
import test_raw_abi_replay as T
from cekirdekler_b200._native import shim
def p(*a): print(*a, flush=True)
T.L = T.Recorder(); L = T.L
w = T.RawWorker(); p("worker ok")
off, glob, loc = w.rng(0), w.rng(T.N), w.rng(T.LOCAL)
w.a[:] = 3.0
L.writeToBuffer(w.q, w.buf_a, w.h_a); L.finish(w.q); p("upload ok")
w.b[:] = -5.0
ue = L.createUserEvent(w.ctx); p("created", ue)
L.addUserEvent(w.q, ue); p("added; last error:", shim().ckGetLastError())
p("compute rc", L.compute(w.q, w.kernels["scaleAdd"], off, glob, loc))
L.readFromBuffer(w.q, w.buf_b, w.h_b); p("read enqueued")
L.flush(w.q); p("flushed")
time.sleep(0.05); p("still gated:", bool((w.b == -5.0).all()))
L.triggerUserEvent(ue); p("triggered")
L.finish(w.q); p("finished:", w.b[:4])
