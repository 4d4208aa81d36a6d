"""Worker of tests/test_gpu_nccl.py (launched by torch.distributed.run, one process per GPU, backend nccl):
rank r encodes its shard of a seeded batch, the encoder-weight gradients are all-reduced through
ggpm_b200.dist.GradBucket, and rank 0 checks that the reduced gradient equals the single-GPU gradient of the
concatenated batch (SURVEY 8(e) / SURVEY 4 item 6).  Prints one JSON line on rank 0."""
import json
import os
import sys

import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))


def main():
    import ggpm_b200
    from ggpm_b200 import ingest, synth, tensorize
    from ggpm_b200.dist import GradBucket
    mode = sys.argv[1] if len(sys.argv) > 1 else "equal"
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    H, depth, B = 250, 20, 32 * world
    torch.manual_seed(0)
    model = ggpm_b200.HierMPNEncoder(synth.SizeOnlyPairVocab(), synth.SizeOnlyAtomVocab(), "LSTM", H, H, depth, depth, 0.0)
    for p in model.parameters():
        torch.nn.init.normal_(p, std=0.05) if p.dim() == 1 else torch.nn.init.xavier_normal_(p)
    model = model.to(dev)
    mols = synth.random_batch(9090, B)
    if mode == "equal":
        mine = mols[rank::world]
    else:                                   # uneven shards: rank 0 holds a quarter of what the others hold
        cut = B // (4 * world)
        bounds = [0, cut] + [cut + (B - cut) * i // (world - 1) for i in range(1, world)]
        mine = mols[bounds[rank]:bounds[rank + 1]]

    def grads_of(batch):
        tree, graph = ingest.pack(tensorize.tensorize(batch)).to_device(dev)
        model.zero_grad(set_to_none=True)
        model(tree, graph)[0].sum().backward()

    early = None
    if mode == "equal":
        # grouped, overlapped reduction: step 0 learns which parameters receive gradients, step 1 launches the early
        # group's all-reduce on the side stream from the autograd hooks while the atom layer's backward still runs
        bucket = GradBucket.for_encoder(model)
        for _ in range(2):
            bucket.arm()
            grads_of(mine)
            early = bool(bucket._launched[0])
            bucket.all_reduce_mean()
    else:
        bucket = GradBucket(model.parameters())
        grads_of(mine)
        bucket.all_reduce_mean(weight=float(len(mine)))
    torch.cuda.synchronize()
    reduced = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
    # every rank must hold the same reduced gradient
    flat = torch.cat([g.reshape(-1) for g in reduced.values()])
    ref = flat.clone()
    dist.broadcast(ref, src=0)
    same = bool(torch.equal(flat, ref))
    ok = torch.tensor([int(same)], device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    worst, worst_key = 0.0, None
    if rank == 0:
        grads_of(mols)                      # single GPU, concatenated batch: loss = sum over molecules
        denom = float(world) if mode == "equal" else None
        for k, p in model.named_parameters():
            full = p.grad
            if full is None:
                assert float(reduced[k].abs().max()) == 0.0, k
                continue
            if mode == "equal":
                want = full / denom
            else:
                # weighted mean of per-rank gradients: sum_r n_r g_r / sum_r n_r; g_r summed over the rank's molecules,
                # so it is not a rescaling of the full gradient — recompute it shard by shard
                want = None
            if want is None:
                continue
            e = float((reduced[k].double() - want.double()).abs().max() / want.double().abs().max().clamp_min(1e-30))
            if e > worst:
                worst, worst_key = e, k
        if mode != "equal":
            acc = {k: torch.zeros_like(p, dtype=torch.float64) for k, p in model.named_parameters()}
            cut = B // (4 * world)
            bounds = [0, cut] + [cut + (B - cut) * i // (world - 1) for i in range(1, world)]
            for r in range(world):
                part = mols[bounds[r]:bounds[r + 1]]
                grads_of(part)
                for k, p in model.named_parameters():
                    if p.grad is not None:
                        acc[k] += len(part) * p.grad.double()
            for k in acc:
                want = acc[k] / B
                if float(want.abs().max()) == 0.0:
                    continue
                e = float((reduced[k].double() - want).abs().max() / want.abs().max())
                if e > worst:
                    worst, worst_key = e, k
        print(json.dumps({"mode": mode, "world": world, "same_on_all_ranks": bool(ok.item()), "worst_rel_err": worst,
                          "worst_key": worst_key, "backend": dist.get_backend(), "early_group_launched_in_backward": early}))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
