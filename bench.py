#!/usr/bin/env python
"""Benchmark of the hierarchical message-passing encoder hot path (BASELINE.json metric:
"HierMPN encoder molecules/sec (fwd+bwd)").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg1|cfg2|cfg3s|cfg3|cfg4|cfg5|cfg5h|cfg2g]

One step = HierMPNEncoder forward + backward (loss = hroot.sum(), the only output the VAEs consume,
ggpm/property_vae.py:50) over one synthetic batch.  The default workload is BASELINE config #4, the one the
metric is quoted on: the synthetic throughput sweep (molecular graphs of ~30 atoms / ~12 motifs, depth 20/20,
LSTM, hidden = embed = 250) streamed in batches of 4096 molecules per GPU; `--workload cfg2` is the encoder of
one vae_train.py pre-training step (pretrained_w_tie_embedding_configs.json, dropout 0.1, batch 256), the other
names are the remaining BASELINE configs.  For N > 1 every rank encodes its own batches (weak scaling, no
communication on the encoder path) and the encoder-weight gradients are all-reduced with NCCL inside the step.

Prints ONE JSON line (rank 0).  `value` is timed with inputs resident in HBM; `e2e` goes through the public
module API from pinned HOST index tensors (H2D copy + loss read-back inside the timed region).
`--impl reference` times the reference's OWN HierMPNEncoder / MotifEncoder (unmodified ggpm/encoder.py, staged
under oracle/_ref/ by oracle/make_ref.py, rdkit stubbed) on the host cores with the same hyper-parameters, dropout
setting, mode and loss; `cpu_baseline.kind` is "reference" then, "port" only if the staged copy is missing (the
oracle restatement is timed instead).
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

# batches differ in size, so the saved-activation blocks (tens of GB each at cfg3) differ from step to step: without
# expandable segments the caching allocator fragments and cfg3 (B = 4096, H = 500) does not fit in 180 GB
os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: hidden, depthT, depthG, dropout, batch per GPU
    "cfg1": dict(hidden=250, depthT=20, depthG=20, dropout=0.0, batch=20,
                 desc="configs/configs.json encoder, 20 synthetic molecules"),
    "cfg2": dict(hidden=250, depthT=20, depthG=20, dropout=0.1, batch=256,
                 desc="encoder of one vae_train.py step, pretrained_w_tie_embedding_configs.json, batch 256"),
    "cfg3s": dict(hidden=500, depthT=20, depthG=20, dropout=0.0, batch=1024,
                  desc="double-hidden encoder (hidden 500), batch 1024 synthetic graphs (profiling-sized cfg3)"),
    "cfg3": dict(hidden=500, depthT=20, depthG=20, dropout=0.0, batch=4096,
                 desc="double-hidden encoder (hidden 500), batch 4096 synthetic graphs"),
    "cfg4": dict(hidden=250, depthT=20, depthG=20, dropout=0.0, batch=4096, pool=16,
                 desc="throughput sweep: synthetic molecular graphs (~30 atoms, ~12 motifs, depth 20) streamed in batches "
                      "of 4096 per GPU; the stream REPLAYS a rotating pool of 16 distinct batches per rank (65 536 "
                      "distinct molecules per GPU, topologies repeat every 16 steps); --steps 245 at 1 GPU (31 at 8) "
                      "covers 1M molecules"),
    "cfg5": dict(hidden=250, depthT=20, depthG=20, dropout=0.1, batch=256, encoder="motif", latent=24,
                 desc="PropertyVAE encoder side of one pre-training step: MotifEncoder + R_mean/R_var rsample/KL "
                      "(ggpm/property_vae.py:79-99), batch 256 per GPU, decoder excluded"),
    "cfg2g": dict(hidden=250, depthT=20, depthG=20, dropout=0.1, batch=256, rnn="GRU",
                  desc="cfg2 with rnn_type GRU (constructible, ggpm/encoder.py:21-22; no shipped config uses it), batch 256"),
    "cfg5h": dict(hidden=250, depthT=20, depthG=20, dropout=0.1, batch=256, latent=24,
                  desc="HierPropertyVAE ('hier-prop', ggpm/opvnet.py:7) encoder side of one pre-training step: HierMPNEncoder + "
                       "R_mean/R_var rsample/KL (ggpm/property_vae.py:48-52), batch 256 per GPU, decoder excluded"),
}
METRIC = "HierMPN encoder molecules/sec (fwd+bwd)"
UNIT = "molecules/s"
PROF_CLASSES = ["gather_sum", "gemm_tcgen05", "gemm_fp32_simt", "gemm_tcgen05_wgrad", "cell_fwd", "cell_bwd", "misc"]


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(hbm=float(p["hbm_gbs"]), bf16=float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), src="measured")
    return dict(hbm=6650.0, bf16=1590.0, src="fallback")


class ClockSampler:
    """SM clock and throttle reasons sampled every 200 ms while the timed region runs, through NVML in-process
    (nvidia_ml_py).  Spawning `nvidia-smi -lms 200` instead was measured to stall kernel launches on multi-GPU hosts
    (3x longer steps while it polled), so it is only the fallback, at a 1 s period."""

    def __init__(self, gpu_index):
        self.gpu, self.rows, self.stop_flag, self.thread, self.proc = gpu_index, [], False, None, None
        self.mode = None
        self.call_ms = []
        self.period = float(os.environ.get("GGPM_BENCH_CLOCK_PERIOD", "0.2"))

    def _nvml_loop(self):
        import pynvml as nv
        R = [("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown")
              else nv.nvmlClocksThrottleReasonHwSlowdown),
             ("hw_thermal_slowdown", getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", None)
              or nv.nvmlClocksThrottleReasonHwThermalSlowdown),
             ("sw_thermal_slowdown", getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", None)
              or nv.nvmlClocksThrottleReasonSwThermalSlowdown),
             ("sw_power_cap", getattr(nv, "nvmlClocksEventReasonSwPowerCap", None) or nv.nvmlClocksThrottleReasonSwPowerCap)]
        self._reasons = R
        while not self.stop_flag:
            self.sample_now()
            time.sleep(self.period)

    def sample_now(self):
        """one synchronous NVML sample (also called by the timing loop right after the last timed step is enqueued,
        while the GPU is still working through the queue)"""
        if self.mode != "nvml":
            return
        import pynvml as nv
        try:
            t0 = time.perf_counter()
            sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
            mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            self.call_ms.append((time.perf_counter() - t0) * 1e3)
            R = getattr(self, "_reasons", None) or []
            self.rows.append((float(sm), float(self.max_sm), [n for n, bit in R if mask & bit]))
        except Exception:
            pass

    def _smi_loop(self):
        for line in self.proc.stdout:
            r = [c.strip() for c in line.split(",")]
            try:
                names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
                self.rows.append((float(r[0]), float(r[1]), [n for n, v in zip(names, r[2:6]) if v.lower().startswith("active")]))
            except Exception:
                continue

    def start(self, threaded=True):
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = self.gpu
            if vis:
                ids = [v for v in vis.split(",") if v.strip() != ""]
                if self.gpu < len(ids) and ids[self.gpu].strip().isdigit():
                    phys = int(ids[self.gpu])
            self.h = None
            try:                                     # exact device even when CUDA_VISIBLE_DEVICES holds UUIDs
                uuid = str(torch.cuda.get_device_properties(self.gpu).uuid)
                self.h = nv.nvmlDeviceGetHandleByUUID(uuid if uuid.startswith("GPU-") else "GPU-" + uuid)
            except Exception:
                self.h = None
            if self.h is None:
                self.h = nv.nvmlDeviceGetHandleByIndex(phys)
            self.max_sm = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
            self.mode = "nvml"
            # the first NVML query of a process takes ~30 ms inside the driver (measured) and would land in the
            # timed region: pay for it here
            nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
            nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            R = [("hw_slowdown", "HwSlowdown"), ("hw_thermal_slowdown", "HwThermalSlowdown"),
                 ("sw_thermal_slowdown", "SwThermalSlowdown"), ("sw_power_cap", "SwPowerCap")]
            self._reasons = [(n, getattr(nv, "nvmlClocksEventReason" + k, None) or getattr(nv, "nvmlClocksThrottleReason" + k))
                             for n, k in R]
            if threaded:
                self.thread = threading.Thread(target=self._nvml_loop, daemon=True)
                self.thread.start()
            return
        except Exception:
            pass
        try:
            q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                 "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "1000"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.mode = "nvidia-smi"
            self.thread = threading.Thread(target=self._smi_loop, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def stop(self):
        if self.mode is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no NVML / nvidia-smi"]}
        time.sleep(0.25)
        self.stop_flag = True
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        if self.thread is not None:
            self.thread.join(timeout=2)
        sm = [r[0] for r in self.rows]
        mx = [r[1] for r in self.rows]
        reasons = sorted({n for r in self.rows for n in r[2]})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "source": self.mode,
                "nvml_call_ms_max": round(max(self.call_ms), 2) if self.call_ms else None}


_MEAN_ROWS = {}


def _mol_rows(m):
    """message rows a molecule contributes to the three layers: atom messages + 2 x tree messages"""
    return len(m.bonds) + 2 * len(m.tree_edges)


def balanced_batch(seed, B, slack=0.125):
    """B molecules whose total message-row count is within ~0.05 % of B x (mean rows per molecule of the generator):
    drawn as B + slack*B candidates, then single swaps between the chosen and the spare ones.  Data-parallel steps are
    synchronous, so the step time is the MAX over ranks: SURVEY 8(e) asks to equalise the sum of messages per rank rather
    than the molecule count.  The target is a constant of the generator (fixed-seed sample), identical on every rank and
    for every N, so no cross-rank exchange is needed and N = 1 runs the same distribution."""
    from ggpm_b200 import synth
    if "mean" not in _MEAN_ROWS:
        sample = synth.random_batch(987654321, 2048)
        _MEAN_ROWS["mean"] = float(np.mean([_mol_rows(m) for m in sample]))
    extra = max(8, int(B * slack))
    mols = synth.random_batch(seed, B + extra)
    w = np.asarray([_mol_rows(m) for m in mols], dtype=np.int64)
    chosen, spare = list(range(B)), list(range(B, B + extra))
    target = int(round(B * _MEAN_ROWS["mean"]))
    for _ in range(64):
        diff = int(w[chosen].sum()) - target                 # > 0: too heavy
        if abs(diff) <= max(1, target // 4000):
            break
        gain = w[chosen][:, None] - w[spare][None, :]        # swapping chosen i for spare j lowers the sum by gain[i, j]
        i, j = np.unravel_index(np.argmin(np.abs(gain - diff)), gain.shape)
        if abs(diff - int(gain[i, j])) >= abs(diff):
            break
        chosen[i], spare[j] = spare[j], chosen[i]
    return [mols[i] for i in chosen]


def make_batches(workload, n_batches, seed, balance=True):
    from ggpm_b200 import synth, tensorize
    out = []
    for i in range(n_batches):
        mols = (balanced_batch(seed * 1000 + i, workload["batch"]) if balance
                else synth.random_batch(seed * 1000 + i, workload["batch"]))
        tree, graph = tensorize.tensorize(mols)      # C++ index builder (bit-identical to synth.tensorize)
        out.append((tree, graph, synth.batch_stats(tree, graph)))
    return out


def algorithmic_flops(stats, H, depthT, depthG, B, motif_only=False):
    """minimal-formulation FLOPs of one HierMPNEncoder forward (LSTM); fwd+bwd = 3x"""
    def layer(E, N, IN, F, D):
        return 2.0 * E * IN * 4 * H + D * 8.0 * E * H * H + 2.0 * N * (F + H) * H
    if motif_only:      # MotifEncoder: one tree layer + W_root (+ the two latent projections, negligible)
        return layer(stats["Et"], stats["Nt"], H + 20, H, depthT) + 2.0 * B * 2 * H * H
    f = layer(stats["Ea"], stats["Na"], 62, 38, depthG) + 2 * layer(stats["Et"], stats["Nt"], H + 20, H, depthT)
    f += 2 * (2.0 * stats["Nt"] * 2 * H * H) + 2.0 * B * 2 * H * H
    return f


# --------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own encoder classes on the host cores
# --------------------------------------------------------------------------------------------------
def _reference_step_fn(workload, sample_mols):
    """-> (step(), kind): one forward + backward of the reference's own HierMPNEncoder / MotifEncoder (kind
    "reference"; unmodified ggpm/encoder.py from oracle/_ref or /root/reference) on a `sample_mols`-molecule batch of
    the workload's generator, same hyper-parameters / dropout / train mode / loss as our arm; the oracle restatement
    (kind "port") only when no copy of the reference is reachable."""
    from ggpm_b200 import synth
    from oracle import ref_loader
    H = workload["hidden"]
    mols = synth.random_batch(424242, sample_mols)
    tree, graph = synth.tensorize(mols)
    as_long = lambda ts: [torch.from_numpy(np.asarray(t)).long() if not isinstance(t, list) else t for t in ts]
    tree, graph = as_long(tree), as_long(graph)
    motif = workload.get("encoder") == "motif"
    L = workload.get("latent", 0)
    if ref_loader.reference_available():
        ref = ref_loader.load_reference()
        ref.nnutils.device = torch.device("cpu")          # the reference pins cuda:0 when it sees a GPU (nnutils.py:9-10)
        torch.manual_seed(0)
        cls = ref.encoder.MotifEncoder if motif else ref.encoder.HierMPNEncoder
        model = cls(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), workload.get("rnn", "LSTM"), H, H, workload["depthT"],
                    workload["depthG"], workload["dropout"])
        if L:
            model.R_mean, model.R_var = torch.nn.Linear(H, L), torch.nn.Linear(H, L)
        for p in model.parameters():                       # vae_train.py:49-53
            if p.dim() == 1:
                torch.nn.init.constant_(p, 0)
            else:
                torch.nn.init.xavier_normal_(p)
        model.train()

        def step():
            model.zero_grad(set_to_none=True)
            if L:          # (Hier)PropertyVAE.forward's encoder side (ggpm/property_vae.py:48-52, 95-99): encoder + rsample / KL
                root = model(tree)[0] if motif else model(tree, graph)[0]
                z_mean = model.R_mean(root)
                z_log_var = -torch.abs(model.R_var(root))
                kl = -0.5 * torch.sum(1.0 + z_log_var - z_mean * z_mean - torch.exp(z_log_var)) / root.size(0)
                z = z_mean + torch.exp(z_log_var / 2) * torch.randn_like(z_mean)
                (z.sum() + kl).backward()
            else:
                model(tree, graph)[0].sum().backward()
        return step, "reference"

    from oracle import encoder_oracle as eo
    sd = {k: v.requires_grad_(True)
          for k, v in eo.random_state("motif" if motif else "hier", workload.get("rnn", "LSTM"), H, H, 0, bias_std=0.0).items()}
    if L:
        from oracle import latent_oracle as lo
        lat = [torch.randn(L, H) / H ** 0.5, torch.zeros(L), torch.randn(L, H) / H ** 0.5, torch.zeros(L)]
        lat = [t.requires_grad_(True) for t in lat]

    def step():
        for v in sd.values():
            v.grad = None
        if L:
            root = (eo.motif_encoder(sd, workload.get("rnn", "LSTM"), workload["depthT"], tree)[0] if motif else
                    eo.hier_encoder(sd, workload.get("rnn", "LSTM"), workload["depthT"], workload["depthG"], tree, graph)[0])
            z, kl = lo.rsample(root, *lat, torch.randn(root.shape[0], L), True)
            (z.sum() + kl).backward()
        else:
            eo.hier_encoder(sd, workload.get("rnn", "LSTM"), workload["depthT"], workload["depthG"], tree, graph)[0].sum().backward()
    return step, "port"


def cpu_reference_rate(workload, warmup, steps, budget_s):
    """times `steps` reference steps after `warmup`; the per-step sample (molecules) is the largest of
    32, 64, 128, 256 (capped by the workload's batch) whose (warmup + steps) steps fit `budget_s` on this box,
    judged from one probe step on 32 molecules.  -> (molecules/s, s/step, sample, kind)"""
    torch.set_num_threads(os.cpu_count() or 1)
    probe, kind = _reference_step_fn(workload, min(32, workload["batch"]))
    probe()                                   # first call pays for thread-pool start-up
    t0 = time.perf_counter()
    probe()
    per_mol = (time.perf_counter() - t0) / min(32, workload["batch"])
    sample = min(32, workload["batch"])
    for cand in (64, 128, 256):
        if cand <= workload["batch"] and per_mol * cand * (warmup + steps) <= budget_s:
            sample = cand
    step, kind = _reference_step_fn(workload, sample)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / max(steps, 1)
    return sample / dt, dt, sample, kind


def _sample_note(workload, sample, kind, dt, cores):
    from oracle import ref_loader
    where = "oracle/_ref staged copy" if ref_loader.reference_is_staged_copy() else ref_loader.REFERENCE_ROOT
    what = ("the reference's own %s (unmodified ggpm/encoder.py, " + where + ")" if kind == "reference" else
            "the oracle restatement of %s (reference's padded PyTorch formulation)") % (
                "MotifEncoder + rsample/KL" if workload.get("encoder") == "motif" else "HierMPNEncoder")
    scale = "" if sample == workload["batch"] else (
        "; the workload's batch is %d, CPU time is linear in molecules, so molecules/s carries over" % workload["batch"])
    return ("%d-molecule batch of the same generator per step, fwd+bwd %.2f s/step, %s, fp32, train mode, dropout %.1f, "
            "%d torch threads%s" % (sample, dt, what, workload["dropout"], cores, scale))


def run_reference(args, workload):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rate, dt, sample, kind = cpu_reference_rate(workload, args.warmup, args.steps, budget_s=150.0)
    cores = os.cpu_count() or 1
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload + ": " + workload["desc"], "hidden": workload["hidden"],
                   "depthT": workload["depthT"], "depthG": workload["depthG"], "rnn_type": workload.get("rnn", "LSTM"),
                   "dropout": workload["dropout"], "batch_per_gpu": workload["batch"]},
        "sample_molecules": sample,
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": _sample_note(workload, sample, kind, dt, cores)},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------
def run_ours(args, workload):
    import torch.distributed as dist
    import ggpm_b200
    from ggpm_b200 import _lib, nnutils, synth
    from ggpm_b200.dist import GradBucket

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the encoder path has no CPU fallback (use --impl reference "
                         "for the CPU oracle)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if args.gpus != world and rank == 0 and world > 1:
        print("warning: --gpus %d but WORLD_SIZE %d" % (args.gpus, world), file=sys.stderr)

    H = workload["hidden"]
    torch.manual_seed(0)
    motif = workload.get("encoder") == "motif"
    enc_cls = ggpm_b200.MotifEncoder if motif else ggpm_b200.HierMPNEncoder
    model = enc_cls(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), workload.get("rnn", "LSTM"), H, H,
                    workload["depthT"], workload["depthG"], workload["dropout"])
    if workload.get("latent"):            # the VAEs keep the latent head next to the encoder (property_vae.py:23-24, 86-87)
        model.R_mean = torch.nn.Linear(H, workload["latent"])
        model.R_var = torch.nn.Linear(H, workload["latent"])
    for p in model.parameters():          # vae_train.py:49-53
        if p.dim() == 1:
            torch.nn.init.constant_(p, 0)
        else:
            torch.nn.init.xavier_normal_(p)
    model = model.to(dev).train()
    # late group = the atom layer; the rest is reduced under its backward (GGPM_BUCKET_OVERLAP=0: one group, reduced after
    # the backward pass — A/B of the overlap)
    if os.environ.get("GGPM_BUCKET_OVERLAP", "1") == "0":
        bucket = GradBucket(model.parameters())
    else:
        bucket = GradBucket.for_encoder(model)

    n_batches = min(4, max(1, args.warmup))      # every batch shape is seen once before the timed region (allocator warm)
    if workload["batch"] >= 4096:
        n_batches = min(n_batches, 2)
    n_batches = workload.get("pool", n_batches)
    if args.pool:
        n_batches = args.pool
    # (diagnostics: GGPM_BENCH_SAME_SEED=1 gives every rank the same batches, GGPM_BENCH_NO_REDUCE=1 skips the collective)
    batches = make_batches(workload, n_batches, seed=1 if os.environ.get("GGPM_BENCH_SAME_SEED") == "1" else rank + 1,
                           balance=not args.no_balance)
    # resident inputs of the device-timed arm: the batch in the ingest format (index tensors + the CSR / reverse CSR the
    # host prepared), already in HBM; --resident make_cuda keeps the reference's int64 tensors instead (index structures
    # are then rebuilt on the device every step)
    from ggpm_b200 import ingest
    packed = [ingest.pack((t, None if motif else g)) for t, g, _ in batches]
    if args.resident == "packed":
        dev_batches = [pb.to_device(dev) for pb in packed]
    else:
        dev_batches = [nnutils.make_cuda((t, g)) for t, g, _ in batches]
    stats = batches[0][2]

    from ggpm_b200 import latent as latent_head

    reduce_grads = world > 1 and os.environ.get("GGPM_BENCH_NO_REDUCE") != "1"

    def forward_loss(tree, graph):
        if workload.get("latent"):     # encoder -> rsample -> (z, KL): the encoder side of (Hier)PropertyVAE.forward
            root = model(tree)[0] if motif else model(tree, graph)[0]
            z, kl = latent_head.rsample(root, model.R_mean, model.R_var, True)
            return z.sum() + kl
        return model(tree, graph)[0].sum()

    def step(tree, graph, read_loss=False):
        model.zero_grad(set_to_none=True)
        loss = forward_loss(tree, graph)
        if reduce_grads:
            bucket.arm()
        loss.backward()
        if reduce_grads:
            bucket.all_reduce_mean()
        return float(loss.item()) if read_loss else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, warmup, steps, end_sampler=None):
        for i in range(warmup):
            fn(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(warmup + i)
        e1.record()
        if end_sampler is not None:      # the host runs ahead of the device: the GPU is still inside the timed steps
            busy = not e1.query()
            end_sampler.sample_now()
            end_sampler.in_region = getattr(end_sampler, "in_region", 0) + int(busy and not e1.query())
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # ---- settle the process before anything is timed (not part of W): the first steps of a fresh process on a fresh
    # box pay for allocator growth, lazy module loading and clock ramp-up (a 2x slower first second was measured);
    # run untimed steps until two consecutive ones agree within 5 % (at most 40)
    # (ranks must run the same number of steps — the step holds a collective — so N > 1 runs a fixed 12)
    prev = None
    for i in range(40 if world == 1 else 12):
        t0 = time.perf_counter()
        step(*dev_batches[i % n_batches])
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world == 1 and prev is not None and i >= 2 * n_batches and abs(dt - prev) <= 0.05 * prev:
            break
        prev = dt
    barrier()

    # ---- device-resident throughput ("value") -------------------------------------------------------
    # Two passes over the same W + K steps.  Pass A runs with the NVML sampler thread (clock / throttle samples every
    # 200 ms under load).  On some boxes of this pool one NVML query holds a driver lock for tens of ms and stalls the
    # kernel launches behind it (steps 1.1-4x longer, measured), so the reported time comes from pass B, which has no
    # sampler thread and takes its clock sample right after the last timed step is enqueued, while the device is still
    # working through the queue.  Both times are printed.
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start(threaded=True)
    ms_sampled = timed(lambda i: step(*dev_batches[i % n_batches]), args.warmup, args.steps)
    if rank == 0:
        time.sleep(0.25)
        sampler.stop_flag = True
        if sampler.thread is not None:
            sampler.thread.join(timeout=2)
    launches0 = _lib.launch_count()
    ms = timed(lambda i: step(*dev_batches[i % n_batches]), args.warmup, args.steps, end_sampler=sampler if rank == 0 else None)
    launches = (_lib.launch_count() - launches0) * args.steps // (args.steps + args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None:
        # `samples`: NVML samples of the sampled pass (200 ms period over its W + K steps under load) plus the one taken
        # right after the last step of the reported pass was enqueued (`samples_inside_timed_region`)
        clocks["samples_inside_timed_region"] = int(getattr(sampler, "in_region", 0))
        clocks["ms_per_step_with_sampler_thread"] = ms_sampled / args.steps
    mols_per_step = workload["batch"] * world
    if args.profile_step:
        # one extra step inside cudaProfilerStart/Stop: `ncu --profile-from-start off` captures exactly one step
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        step(*dev_batches[0])
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    value = mols_per_step * args.steps / (ms / 1e3)
    if args.quick:
        if rank == 0:
            print(json.dumps({"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                              "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                              "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "quick": True,
                              "config": {"workload": args.workload + ": " + workload["desc"], "batch_per_gpu": workload["batch"],
                                         "parallelism": "dp%d" % world,
                                         "bucket_overlap": os.environ.get("GGPM_BUCKET_OVERLAP", "1")},
                              "gpu_launches": int(launches), "clocks": clocks}))
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- end to end from pinned host buffers ("e2e") ------------------------------------------------
    # (1) through the ingest API (ggpm_b200.ingest.PackedBatch): the batch was packed once at
    #     preprocess time into one pinned int32 buffer; the step does ONE H2D copy, the encoder, a loss read-back;
    # (2) through the reference's own make_cuda flow (ggpm/nnutils.py:210-214): one copy + one widening kernel per
    #     tensor, index structures built on the device.
    h2d = packed[0].nbytes

    # pipelined the way a training loop runs it: ingest.DeviceLoader copies batch i+1 on a side stream while batch i is
    # encoded, and the loss of step i goes to pinned host memory asynchronously and is read once step i+1 has been
    # enqueued.  Every step still moves its own inputs H2D and its own result D2H inside the timed region.
    n_e2e = min(args.warmup, 3) + args.steps

    def pipelined_pass():
        loader = ingest.DeviceLoader((packed[i % n_batches] for i in range(n_e2e)), dev, depth=2)
        loss_host = torch.empty(n_e2e, dtype=torch.float32).pin_memory()
        loss_ev = [torch.cuda.Event() for _ in range(n_e2e)]
        loss_seen = []

        def e2e_step(i):
            tree, graph = next(loader)
            model.zero_grad(set_to_none=True)
            loss = forward_loss(tree, graph)
            if world > 1:
                bucket.arm()
            loss.backward()
            if world > 1:
                bucket.all_reduce_mean()
            loss_host[i].copy_(loss.detach(), non_blocking=True)
            loss_ev[i].record()
            if i > 0:
                loss_ev[i - 1].synchronize()
                loss_seen.append(float(loss_host[i - 1]))
        ms_ = timed(e2e_step, min(args.warmup, 3), args.steps)
        loss_ev[n_e2e - 1].synchronize()
        loss_seen.append(float(loss_host[n_e2e - 1]))
        assert len(loss_seen) == n_e2e and all(np.isfinite(loss_seen))
        return ms_

    # Each end-to-end leg is timed three times (W + K steps each) and the MEDIAN pass is reported, all three are
    # printed: on some boxes of this pool a single host stall of 50-100 ms (measured: one pass at 21.8 ms/step between
    # passes at 12.9 ms/step, same build) lands inside a 10-step region.
    def median_of_three(run):
        passes = [run() for _ in range(3)]
        return sorted(passes)[1], [p / args.steps for p in passes]

    e2e_ms, e2e_passes = median_of_three(pipelined_pass)
    e2e_value = mols_per_step * args.steps / (e2e_ms / 1e3)

    # the same without any pipelining: copy, encode, loss.item() strictly one after the other
    def e2e_sync_step(i):
        tree, graph = packed[i % n_batches].to_device(dev)
        step(tree, graph, read_loss=True)
    e2e_sync_ms, e2e_sync_passes = median_of_three(lambda: timed(e2e_sync_step, min(args.warmup, 3), args.steps))
    e2e_sync_value = mols_per_step * args.steps / (e2e_sync_ms / 1e3)

    def pin(arrs):
        return [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() if not isinstance(a, list) else a for a in arrs]
    host_batches = [(pin(t), pin(g)) for t, g, _ in batches]
    h2d_mc = sum(a.numel() * a.element_size() for a in host_batches[0][0][:-1] + host_batches[0][1][:-1])

    def e2e_mc_step(i):
        t, g = host_batches[i % n_batches]
        tree = [a.to(dev, non_blocking=True).long() for a in t[:-1]] + [t[-1]]
        graph = [a.to(dev, non_blocking=True).long() for a in g[:-1]] + [g[-1]]
        step(tree, graph, read_loss=True)
    e2e_mc_ms, e2e_mc_passes = median_of_three(lambda: timed(e2e_mc_step, min(args.warmup, 3), args.steps))
    e2e_mc_value = mols_per_step * args.steps / (e2e_mc_ms / 1e3)

    # ---- per-kernel-class timing for the roofline (separate instrumented pass, rank 0) --------------
    roofline, kernels, rooflines = None, None, None
    peaks = load_peaks()
    traffic = {}
    try:      # HBM bytes per launch from the committed ncu --set full capture (atom-layer launch of a cfg2 step)
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
        if tj.get("workload") == args.workload:
            traffic = tj.get("per_launch_bytes", {})
        else:   # captures of other workloads: {"others": {"cfg4": {"per_launch_bytes": {...}, "source": "..."}}}
            traffic = tj.get("others", {}).get(args.workload, {}).get("per_launch_bytes", {})
    except Exception:
        traffic = {}
    if rank == 0:
        lib = _lib.backend()
        _lib.call("ggpm_prof_enable", 1)
        prof_steps = 2
        for i in range(prof_steps):
            model.zero_grad(set_to_none=True)
            forward_loss(*dev_batches[i % n_batches]).backward()
        n = len(PROF_CLASSES)
        ms_c, cnt_c, work_c = (ctypes.c_double * n)(), (ctypes.c_longlong * n)(), (ctypes.c_double * n)()
        _lib.call("ggpm_prof_collect", ctypes.cast(ms_c, ctypes.c_void_p), ctypes.cast(cnt_c, ctypes.c_void_p),
                  ctypes.cast(work_c, ctypes.c_void_p), n)
        _lib.call("ggpm_prof_enable", 0)
        total = sum(ms_c) or 1.0
        kernels = {PROF_CLASSES[c]: {"ms_per_step": ms_c[c] / prof_steps, "launches_per_step": cnt_c[c] // prof_steps,
                                     "share": ms_c[c] / total} for c in range(n) if cnt_c[c]}
        gemm_classes = [c for c in (1, 2, 3) if cnt_c[c]]
        # HBM-side classes: ALGORITHMIC bytes per step, per class, from SURVEY 8(d) on the unpadded width H (fp32).
        # Per LSTM round and layer with E messages and nnz true neighbour entries, one row = H * 4 bytes:
        #   gather class   forward: the neighbour gather of h (nnz rows read) + S written (E rows: the large-layer path
        #                  forms S with a gather kernel); backward mirrors it over the reverse CSR (nnz + E);
        #                  plus the node / cluster aggregations (nnz_a + N, nnz_c + Nt rows) each way
        #   cell_fwd       neighbour rows of c and of the forget projection U_f h (2 nnz read), h', c' written (2 E)
        #   cell_bwd       the same two neighbour tables re-read for the forget gate (2 nnz), the two scatter targets
        #                  dC_prev / dP (2 nnz), dH, dC and the saved c_t read (3 E)
        #   + 4 bytes per index entry (nnz + E) per pass.  Everything else a kernel moves (gate pre-activations, the
        #   x-projection rows, dX, fp16 planes) is implementation traffic and counts against `frac`.
        rowb = H * 4.0
        layers = ((stats["Ea"], stats["nnz_bA"], workload["depthG"], stats["Na"], stats["nnz_aA"]),
                  (stats["Et"], stats["nnz_bT"], workload["depthT"], stats["Nt"], stats["nnz_aT"]),
                  (stats["Et"], stats["nnz_bT"], workload["depthT"], stats["Nt"], stats["nnz_aT"]))
        by_cls = {0: 0.0, 4: 0.0, 5: 0.0}
        for E, nnz, D, N, nnz_a in (layers[1:2] if motif else layers):
            idx = 4.0 * (nnz + E)
            # layers below the library's 32 768-row threshold gather h inside the forward GEMM's operand prologue
            # (no gather-class launch forward): only the backward mirror is a gather kernel there
            passes = 2 if (E + 1) >= 32768 else 1
            by_cls[0] += passes * (D - 1) * ((nnz + E) * rowb + idx) + 2 * ((nnz_a + N) * rowb + 4.0 * (nnz_a + N))
            by_cls[4] += (D - 1) * ((2.0 * nnz + 2.0 * E) * rowb + idx) + 2.0 * E * rowb
            by_cls[5] += (D - 1) * ((4.0 * nnz + 3.0 * E) * rowb + idx) + 3.0 * E * rowb
        if not motif:
            by_cls[0] += 2 * ((stats["nnz_c"] + stats["Nt"]) * rowb + 4.0 * (stats["nnz_c"] + stats["Nt"]))
        flop_scale = (H / float((H + 31) // 32 * 32)) ** 2     # the library counts 2 M N K on the padded width Hp

        def class_roofline(c):
            name = PROF_CLASSES[c]
            if c in gemm_classes:
                tf = work_c[c] * flop_scale / (ms_c[c] * 1e-3) / 1e12
                return {"kernel": name, "bound": "tensor", "achieved": tf, "peak": peaks["bf16"], "unit": "TFLOP/s",
                        "frac": tf / peaks["bf16"], "traffic": traffic.get(name),
                        "flops_per_launch": work_c[c] * flop_scale / cnt_c[c], "avg_launch_ms": ms_c[c] / cnt_c[c],
                        "executed_tensor_frac": 3.0 * work_c[c] / (ms_c[c] * 1e-3) / 1e12 / peaks["bf16"],
                        "peak_kind": "dense bf16 cuBLAS, sustained, of " + peaks["src"] +
                                     "; algorithmic FLOPs on the unpadded H; f16x3 executes 3 tensor FLOPs per "
                                     "algorithmic FLOP on the padded Hp (executed_tensor_frac)"}
            gbs = by_cls.get(c, 0.0) * prof_steps / (ms_c[c] * 1e-3) / 1e9
            return {"kernel": name, "bound": "hbm", "achieved": gbs, "peak": peaks["hbm"], "unit": "GB/s",
                    "frac": gbs / peaks["hbm"], "traffic": traffic.get(name), "avg_launch_ms": ms_c[c] / cnt_c[c],
                    "algorithmic_bytes_per_launch": by_cls.get(c, 0.0) / max(cnt_c[c] // prof_steps, 1),
                    "peak_kind": "copy bandwidth, of " + peaks["src"] + "; per-class algorithmic bytes (bench.py)"}

        top = max(range(n), key=lambda c: ms_c[c])
        roofline = class_roofline(top)
        # the other large classes, same definitions (the top class flips between cell_bwd and the gate GEMMs run to run)
        rooflines = {PROF_CLASSES[c]: class_roofline(c) for c in (0, 1, 3, 4, 5) if cnt_c[c] and c != top}

    # ---- CPU baseline on the host cores (rank 0, N = 1 only) ----------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        rate, dt, sample, kind = cpu_reference_rate(workload, 1, 2, budget_s=25.0)
        cores = os.cpu_count() or 1
        cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": "1 warm-up + 2 timed steps: " + _sample_note(workload, sample, kind, dt, cores)}

    if rank == 0:
        flops = 3.0 * algorithmic_flops(stats, H, workload["depthT"], workload["depthG"], workload["batch"], motif)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload + ": " + workload["desc"], "hidden": H, "depthT": workload["depthT"],
                       "depthG": workload["depthG"], "rnn_type": workload.get("rnn", "LSTM"), "dropout": workload["dropout"],
                       "batch_per_gpu": workload["batch"], "parallelism": "dp%d" % world,
                       "gemm_engine": {0: "auto (tcgen05 f16x3 + 3xtf32)", 1: "fp32-simt", 2: "tcgen05-3xtf32", 3: "tcgen05-f16x3"}[lib.ggpm_get_gemm_engine()],
                       "l2": "no explicit flush: %d distinct batches rotate and each step streams its saved "
                             "activations (>> 126 MB L2) between forward and backward" % n_batches,
                       "load_balance": ("none (i.i.d. batches)" if args.no_balance else
                                        "every batch is drawn with 12.5 % spare molecules and trimmed by single swaps to the "
                                        "generator's mean message-row count x batch (+-0.05 %), the same constant on every "
                                        "rank and for every N: synchronous data-parallel steps wait for the slowest rank"),
                       "shape_per_batch": stats},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4,
                    "ms_per_step": e2e_ms / args.steps, "passes_ms_per_step": e2e_passes,
                    "api": "ingest.DeviceLoader (one pinned int32 buffer per batch, packed ahead of time, H2D of batch "
                           "i+1 on a side stream) -> HierMPNEncoder.forward -> backward -> async loss read-back, "
                           "read one step later"},
            "e2e_unpipelined": {"value": e2e_sync_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                                "d2h_bytes_per_step": 4, "ms_per_step": e2e_sync_ms / args.steps,
                                "passes_ms_per_step": e2e_sync_passes,
                                "api": "PackedBatch.to_device -> forward -> backward -> loss.item(), no overlap"},
            "e2e_make_cuda": {"value": e2e_mc_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d_mc),
                              "d2h_bytes_per_step": 4, "ms_per_step": e2e_mc_ms / args.steps,
                              "passes_ms_per_step": e2e_mc_passes,
                              "api": "reference flow: per-tensor pinned copy + .long() (make_cuda), index structures "
                                     "built on the device"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roofline,
            "rooflines_other": rooflines,
            "kernels": kernels,
            "algorithmic_tflops": flops * world / (ms / args.steps * 1e-3) / 1e12,
            "cpu_baseline": cpu,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg4", choices=sorted(WORKLOADS))
    ap.add_argument("--engine", type=int, default=None, help="0 auto, 1 fp32 CUDA cores, 2 tcgen05 3xTF32")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--resident", default="packed", choices=["packed", "make_cuda"],
                    help="format of the HBM-resident inputs of the device-timed arm")
    ap.add_argument("--no-balance", action="store_true",
                    help="plain i.i.d. batches instead of batches equalised in total message rows (SURVEY 8(e))")
    ap.add_argument("--quick", action="store_true", help="device-timed value only (scaling sweeps): no e2e legs, no per-class pass")
    ap.add_argument("--pool", type=int, default=0, help="number of distinct batches per rank (profiling runs: 2 starts faster)")
    ap.add_argument("--profile-step", action="store_true",
                    help="run one extra step between cudaProfilerStart/Stop (for ncu --profile-from-start off)")
    args = ap.parse_args()
    workload = WORKLOADS[args.workload]
    if os.environ.get("GGPM_ALLOW_TEST_BACKEND"):
        raise SystemExit("bench.py: GGPM_ALLOW_TEST_BACKEND is set — the C-ABI emulator seam of tests/ must be off "
                         "for any measurement")
    if args.impl == "reference":
        run_reference(args, workload)
        return
    if args.engine is not None:
        from ggpm_b200 import _lib
        _lib.set_gemm_engine(args.engine)
    run_ours(args, workload)


if __name__ == "__main__":
    main()
