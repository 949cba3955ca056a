"""Data-parallel host logic on CPU (gloo, world_size 2): the flat-bucket gradient averaging used by CGANTrainer and
the sharding contract of SURVEY.md §8e: averaged per-shard critic / generator gradients == gradients of the whole batch
(losses are batch means, InstanceNorm / GP / dropout / EIoU are per-sample).  The per-rank compute is the CPU oracle."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import cgan_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, B, S, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from b200cgan.train_step import allreduce_mean_
    torch.set_num_threads(2)
    st = O.init_state(42)
    batch, masks, alphas = O.make_batch(B, S, 0)
    lo, hi = rank * B // world, (rank + 1) * B // world
    sl = lambda t: t[lo:hi]
    d = {k: v.clone() for k, v in st["D"].items()}
    _, dg = O.critic_grads(d, sl(batch["pred"]), sl(batch["gt"]), sl(batch["refined"]), sl(alphas[0]), 1.0)
    _, gg, _ = O.generator_grads(st["G"], {k: v.clone() for k, v in st["D"].items()}, sl(batch["pred"]), sl(batch["refined"]),
                                 sl(batch["pred_box"]), sl(batch["delta_true"]), [sl(m) for m in masks[0]])
    dl, gl = [dg[k] for k in sorted(dg)], [gg[k].clone() for k in sorted(gg)]
    allreduce_mean_(dl, None, world)
    allreduce_mean_(gl, None, world)
    if rank == 0:
        q.put(({k: t.numpy() for k, t in zip(sorted(dg), dl)}, {k: t.numpy() for k, t in zip(sorted(gg), gl)},
               {k: d[k].numpy() for k in d if k.endswith("_u") or k.endswith("_v")}))  # numpy: pickled by value
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_gradients_average_to_full_batch_gradients():
    B, S, world = 4, 32, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, B, S, q)) for r in range(world)]
    for p in procs:
        p.start()
    dgr, ggr, uv = ({k: torch.from_numpy(v) for k, v in m.items()} for m in q.get(timeout=600))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    st = O.init_state(42)
    batch, masks, alphas = O.make_batch(B, S, 0)
    d = {k: v.clone() for k, v in st["D"].items()}
    _, dg = O.critic_grads(d, batch["pred"], batch["gt"], batch["refined"], alphas[0], 1.0)
    _, gg, _ = O.generator_grads(st["G"], {k: v.clone() for k, v in st["D"].items()}, batch["pred"], batch["refined"],
                                 batch["pred_box"], batch["delta_true"], masks[0])
    rel = lambda a, b: float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))
    for k in dg:
        if k in O.D_BIAS_UNOBSERVABLE:
            continue
        assert rel(dgr[k], dg[k]) < 2e-4, (k, rel(dgr[k], dg[k]))
    for k in gg:
        assert rel(ggr[k], gg[k]) < 2e-4, (k, rel(ggr[k], gg[k]))
    for k, v in uv.items():  # spectral-norm buffers depend only on the replicated weights: identical on every rank
        assert torch.equal(v, d[k]), k
