"""Time the incremental encoders over a full teacher-forced decode loop: forward through every
step + backward from seeded cotangents on the rows the decoder reads.  Prints one JSON line: molecules/s on the GPU
(CUDA events, warm-up, median of N), launches per loop, per-class kernel time, and the CPU oracle (the reference's
formulation, torch fp32, all host threads) on a bounded sample of the same workload.

    python tools/inc_time.py [--batch 256] [--hidden 250] [--rnn LSTM] [--iters 5] [--cpu-batch 16]
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from ggpm_b200 import _lib, decode_loop, synth                 # noqa: E402
from ggpm_b200 import encoder as genc                          # noqa: E402


def make_case(args, batch):
    return dict(name="t", kind=args.kind, rnn=args.rnn, hidden=args.hidden, depthT=args.diterT, depthG=args.diterG,
                batch=batch, data_seed=5, weight_seed=6, vocab=(721, 6215), mean_motifs=12.0)


def inputs(case, device):
    import numpy as np
    mols = synth.random_batch(case["data_seed"], case["batch"])
    tree_np, graph_np = synth.tensorize(mols)
    sched = decode_loop.decode_schedule(mols, tree_np, graph_np)
    as_long = lambda ts: [torch.from_numpy(np.ascontiguousarray(t)).long().to(device) for t in ts[:-1]] + [ts[-1]]
    init = torch.randn(case["batch"], case["hidden"], generator=torch.Generator().manual_seed(1)) * 0.5
    return as_long(tree_np), as_long(graph_np), sched, init


def loop(model, hier, tree_t, graph_t, sched, init, device, cots=None):
    outs = decode_loop.run_decode(model, hier, tree_t, graph_t, sched, init, device=device)
    if cots is None:
        gen = torch.Generator().manual_seed(2)
        cots = [torch.randn(tuple(o.shape), generator=gen).to(device) for o in outs]
    sum((o * c).sum() for o, c in zip(outs, cots)).backward()
    return cots


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--hidden", type=int, default=250)
    ap.add_argument("--rnn", default="LSTM")
    ap.add_argument("--kind", default="hier", choices=["hier", "motif"])
    ap.add_argument("--diterT", type=int, default=1)
    ap.add_argument("--diterG", type=int, default=5)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--cpu-batch", type=int, default=16)
    args = ap.parse_args()
    dev = "cuda"
    hier = args.kind == "hier"
    from oracle.ref_loader import FakePairVocab, IdAtomVocab       # vocabulary stubs only
    case = make_case(args, args.batch)
    cls = genc.IncHierMPNEncoder if hier else genc.IncEncoder
    torch.manual_seed(0)
    model = cls(FakePairVocab(), IdAtomVocab(), args.rnn, args.hidden, args.hidden, args.diterT, args.diterG, 0.0).to(dev)
    tree_t, graph_t, sched, init = inputs(case, dev)
    init = init.to(dev).requires_grad_(True)
    cots = loop(model, hier, tree_t, graph_t, sched, init, dev)
    loop(model, hier, tree_t, graph_t, sched, init, dev, cots)
    torch.cuda.synchronize()
    n0 = _lib.launch_count()
    times = []
    for _ in range(args.iters):
        model.zero_grad(set_to_none=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        loop(model, hier, tree_t, graph_t, sched, init, dev, cots)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    launches = (_lib.launch_count() - n0) // args.iters
    ms = sorted(times)[len(times) // 2]
    out = dict(metric="incremental encoder molecules/s (decode loop fwd+bwd)", value=args.batch / ms * 1e3,
               ms_per_loop=ms, steps_per_loop=len(sched["steps"]), library_launches_per_loop=launches,
               config=dict(batch=args.batch, hidden=args.hidden, rnn=args.rnn, kind=args.kind, diterT=args.diterT,
                           diterG=args.diterG))
    # per-class kernel time of one loop (CUDA events around the library's launches)
    import ctypes
    _lib.call("ggpm_prof_enable", 1)
    model.zero_grad(set_to_none=True)
    loop(model, hier, tree_t, graph_t, sched, init, dev, cots)
    torch.cuda.synchronize()
    ms_c = (ctypes.c_double * 7)(); cnt = (ctypes.c_longlong * 7)(); work = (ctypes.c_double * 7)()
    _lib.call("ggpm_prof_collect", ctypes.cast(ms_c, ctypes.c_void_p), ctypes.cast(cnt, ctypes.c_void_p),
              ctypes.cast(work, ctypes.c_void_p), 7)
    _lib.call("ggpm_prof_enable", 0)
    names = ["gather", "gemm_tc", "gemm_simt", "gemm_wgrad", "cell_fwd", "cell_bwd", "misc"]
    out["kernel_ms"] = {n: round(ms_c[i], 3) for i, n in enumerate(names) if cnt[i]}
    out["kernel_launches"] = {n: int(cnt[i]) for i, n in enumerate(names) if cnt[i]}
    # HBM-side class of this path (row gather / scatter / scatter-add): algorithmic bytes as counted by the library
    # (rows_gather (K + 1) M W 4, rows_scatter 2 M W 4, rows_scatter_add 3 M K W 4) over its summed launch time
    peak = 6539.9
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    if cnt[0]:
        ach = work[0] / (ms_c[0] * 1e-3) / 1e9
        out["roofline"] = dict(kernel="rows_gather / rows_scatter / rows_scatter_add", bound="hbm", achieved=round(ach, 1),
                               peak=peak, unit="GB/s", frac=round(ach / peak, 4), traffic=None,
                               avg_launch_us=round(ms_c[0] * 1e3 / cnt[0], 2),
                               note="launches of a few microseconds on tables that fit the L2: latency bound, not HBM bound")

    if args.cpu_batch > 0:
        from oracle import inc_oracle as io, make_golden_inc as mg
        ccase = make_case(args, args.cpu_batch)
        ctree, cgraph, csched, cinit = inputs(ccase, "cpu")
        sd = {k: v.detach().cpu().requires_grad_(True) for k, v in model.state_dict().items()}
        orc = io.IncHierOracle(sd, args.rnn, args.diterT, args.diterG) if hier else io.IncMotifOracle(sd, args.rnn, args.diterT)
        t0 = time.perf_counter()
        o = decode_loop.run_decode(orc, hier, ctree, cgraph, csched, cinit.requires_grad_(True))
        sum(x.sum() for x in o).backward()
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = dict(value=args.cpu_batch / dt, unit="molecules/s", cores=torch.get_num_threads(), kind="port",
                                   sample="%d molecules, one decode loop fwd+bwd, fp32" % args.cpu_batch)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
