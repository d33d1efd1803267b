"""Run the C host's library face (host/libtransit_host.so: transit_init + do_transit [+ run_transit ...]) on a
cfg and save what the device computed as .npz -- spectrum, extinction, computed[], tau, last[], CIA, intensities,
the per-line index trace of chosen layers -- through the host's own context (th_context()) and the C-ABI taps.
Executed as a subprocess by tests/test_fullsize_parity.py (the host has one global state and exits on errors,
like the reference it mirrors).

    python tests/hostrun.py <cfg> <out.npz> [--trace r0,r1,..] [--rerun file.f64 [--batch]] [--no-taps] [-- extra transit args]
"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from transit_b200 import api  # noqa: E402


def main():
    cfg, outp = sys.argv[1], sys.argv[2]
    rest = sys.argv[3:]
    extra = []
    if "--" in rest:
        k = rest.index("--")
        rest, extra = rest[:k], rest[k + 1:]
    trace = [int(x) for x in rest[rest.index("--trace") + 1].split(",")] if "--trace" in rest else []
    rerun = rest[rest.index("--rerun") + 1] if "--rerun" in rest else None
    taps = "--no-taps" not in rest
    L = C.CDLL(os.path.join(ROOT, "host", "libtransit_host.so"))
    argv = [b"transit", b"-c", cfg.encode()] + [a.encode() for a in extra]
    arr = (C.c_char_p * len(argv))(*argv)
    so = os.dup(1)
    os.dup2(os.open(os.devnull, os.O_WRONLY), 1)
    t0 = time.perf_counter()
    L.transit_init(len(argv), arr)
    init_s = time.perf_counter() - t0
    L.get_no_samples.restype = C.c_int
    n = L.get_no_samples()
    P = C.POINTER(C.c_double)
    out = np.zeros(n)
    L.th_context.restype = C.c_void_p
    if trace:                 # the fused line sum leaves its per-line records only when asked
        api.load_library().tb_set_trace(C.c_void_p(L.th_context()), 1)
    if "--fuse-all" in rest:  # every state through the fused kernel, carried Doppler indices by k_lastq_scan
        api.load_library().tb_set_fused(C.c_void_p(L.th_context()), 2)
        api.load_library().tb_set_fused_reach(C.c_void_p(L.th_context()), 4096, 1)
    t0 = time.perf_counter()
    L.do_transit(out.ctypes.data_as(P))
    run_s = time.perf_counter() - t0
    os.dup2(so, 1)
    sz = (C.c_int64 * 6)()
    L.th_sizes(sz)
    nlayer, nmol, niso, nwn, nlines, nang = (int(x) for x in sz)
    L.th_context.restype = C.c_void_p
    d = dict(out=out.copy(), init_s=init_s, run_s=run_s, sizes=np.array([nlayer, nmol, niso, nwn, nlines, nang]))
    if taps:
        ctx = api.Context.adopt(L.th_context(), nwn, nlayer, nmol, niso, nlines, nang)
        d["ext"] = ctx.get_extinction()
        d["computed"] = ctx.get_computed()
        d["tau"], d["last"] = ctx.get_tau()
        d["e_cs"] = ctx.get_cia()
        if nang:
            d["intens"] = ctx.get_intensity()
        for r in trace:
            t = ctx.line_trace(r)
            for k, v in t.items():
                d[f"idx{r}_{k}"] = v
            d[f"idx{r}_ilor"] = ctx.state_ilor(r)
        if trace:                              # profsize[nDop][nLor] of OUR table, for minj / maxj (extinction.c:486-496)
            keys = dict(l.split()[:2] for l in open(cfg) if len(l.split()) >= 2 and not l.startswith("#"))
            ctx.nDop, ctx.nLor = int(keys.get("ndop", 60)), int(keys.get("nlor", 60))
            d["profsize"] = ctx.get_profiles()["psize"]
    if rerun:
        inp = np.fromfile(rerun, "<f8")
        nin = nlayer * (1 + nmol)
        k = inp.size // nin
        outs = np.zeros((k, n))
        os.dup2(os.open(os.devnull, os.O_WRONLY), 1)
        t0 = time.perf_counter()
        for i in range(k):
            a = np.ascontiguousarray(inp[i * nin:(i + 1) * nin])
            L.run_transit(a.ctypes.data_as(P), nin, outs[i].ctypes.data_as(P), n)
        d["rerun_s"] = time.perf_counter() - t0
        os.dup2(so, 1)
        d["rerun_out_all"] = outs
        if "--batch" in rest:          # the same inputs through ONE run_transit_batch call
            b = np.zeros((k, n))
            x = np.ascontiguousarray(inp[:k * nin])
            os.dup2(os.open(os.devnull, os.O_WRONLY), 1)
            t0 = time.perf_counter()
            L.run_transit_batch(x.ctypes.data_as(P), k, b.ctypes.data_as(P))
            d["batch_s"] = time.perf_counter() - t0
            os.dup2(so, 1)
            d["batch_out_all"] = b
            # device milliseconds of tb_run_transit_batch: the timer exists only if the batched path really ran
            ms = C.c_float(-1.0)
            rc = api.load_library().tb_get_kernel_ms(C.c_void_p(L.th_context()), b"batch", C.byref(ms))
            d["batch_device_ms"] = float(ms.value) if rc == 0 else -1.0
    np.savez(outp, **d)
    L.free_memory()


if __name__ == "__main__":
    main()
