#!/usr/bin/env python3
"""DiLoCo training of a GPT-2 style model on synthetic tokens (there is no dataset in this sandbox).

Counterpart of the reference's example scripts python/examples/nanogpt_diloco/{sync_diloco,sync_diloco_concat,
async_diloco}.py: every peer trains H inner AdamW steps, then the peers average their pseudo-gradients and apply an
outer Nesterov step.  Like them it reads a JSON config (--config, sync_diloco.py:160), evaluates every `eval_interval`
iterations, writes a checkpoint then (sync_diloco.py:445-457) and resumes from one with --resume (`init_from='resume'`,
sync_diloco.py:247-262) -- a resumed peer brings its shared-state revision along, which the master accepts as the
group's starting revision.  Differences: parameters and the whole outer step stay on the GPU; peers on one NVLink
domain reduce with the fused kernels; the asynchronous mode overlaps the reduction with the next inner phase on a CUDA
stream instead of a thread, synchronises shared state only when the topology changed, and inserts late joiners with the
reference's second synchronisation (SEND_ONLY / RECEIVE_ONLY, async_diloco.py:551-580).

  # terminal 1 (or pass --spawn-master on rank 0)
  python -m pccl_b200.master --listen-address 0.0.0.0:48148
  # one process per GPU / host
  python examples/diloco.py --master 127.0.0.1:48148 --device cuda:0 [--mode async] [--config examples/configs/tiny.json]
Peers may be started and killed at any time.
"""
import argparse
import json
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import pccl_b200 as pccl  # noqa: E402
from pccl_b200.models import GPT, GPTConfig, flatten_parameters  # noqa: E402
from pccl_b200.parallel import AsyncDiLoCo, DiLoCo  # noqa: E402
from pccl_b200.utils import Profiler  # noqa: E402

DEFAULTS = {"model": "tiny", "inner_steps": 8, "outer_steps": 5, "batch_size": 4, "learning_rate": 3e-4, "outer_learning_rate": 0.7,
            "outer_momentum": 0.9, "eval_interval": 0, "eval_iters": 2, "checkpoint_dir": "", "quant": "none"}


def load_config(args) -> dict:
    cfg = dict(DEFAULTS)
    if args.config:
        cfg.update(json.loads(Path(args.config).read_text()))
    for key in ("model", "inner_steps", "outer_steps", "batch_size", "eval_interval", "checkpoint_dir", "quant"):
        v = getattr(args, key)
        if v is not None:
            cfg[key] = v
    return cfg


@torch.no_grad()
def estimate_loss(model, cfg, model_cfg, dev, gen) -> float:
    model.eval()
    total = 0.0
    for _ in range(cfg["eval_iters"]):
        x = torch.randint(0, model_cfg.vocab_size, (cfg["batch_size"], model_cfg.block_size), generator=gen).to(dev)
        total += float(model(x, x)[1])
    model.train()
    return total / cfg["eval_iters"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--master", default="127.0.0.1:48148")
    ap.add_argument("--spawn-master", action="store_true")
    ap.add_argument("--device", default="cuda:0" if torch.cuda.is_available() else "cpu")
    ap.add_argument("--mode", default="sync", choices=["sync", "async"])
    ap.add_argument("--config", default="", help="JSON file; command-line values override it")
    ap.add_argument("--resume", action="store_true", help="start from checkpoint_dir/ckpt.pt")
    ap.add_argument("--min-peers", type=int, default=1)
    ap.add_argument("--model", default=None, choices=["tiny", "gpt2"])
    ap.add_argument("--inner-steps", dest="inner_steps", type=int, default=None)
    ap.add_argument("--outer-steps", dest="outer_steps", type=int, default=None)
    ap.add_argument("--batch", dest="batch_size", type=int, default=None)
    ap.add_argument("--eval-interval", dest="eval_interval", type=int, default=None)
    ap.add_argument("--checkpoint-dir", dest="checkpoint_dir", default=None)
    ap.add_argument("--quant", default=None, choices=["none", "u8", "fp8"])
    ap.add_argument("--count-from-world", dest="count_from_world", type=int, default=0,
                    help="keep training, but start counting the outer steps only once the run has had this many peers")
    ap.add_argument("--timeline", default="", help="write the section timeline here (.png/.svg) and a Chrome trace next to it")
    args = ap.parse_args()
    cfg = load_config(args)

    dev = torch.device(args.device)
    if dev.type == "cuda":
        torch.cuda.set_device(dev)
    master = None
    if args.spawn_master:
        master = pccl.MasterNode(args.master)
        master.run()
    comm = pccl.Communicator(args.master)
    comm.connect(n_attempts=30)
    while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < args.min_peers:
        comm.update_topology()

    torch.manual_seed(0)  # identical initial weights everywhere; late joiners get them via shared state
    model_cfg = GPTConfig.tiny() if cfg["model"] == "tiny" else GPTConfig.gpt2_124m()
    model = GPT(model_cfg).to(dev)
    inner_flat = flatten_parameters(model)           # the model trains on views into this buffer
    outer = inner_flat.clone()                       # outer ("global") parameters
    quant = {"none": None, "u8": pccl.QuantizationOptions(pccl.DataType.UINT8, pccl.QuantizationAlgorithm.MIN_MAX),
             "fp8": pccl.QuantizationOptions(pccl.DataType.FLOAT8_E4M3, pccl.QuantizationAlgorithm.BLOCK_FP8)}[cfg["quant"]]
    algo = (AsyncDiLoCo if args.mode == "async" else DiLoCo)(comm, outer, lr=cfg["outer_learning_rate"], momentum=cfg["outer_momentum"],
                                                              quantization=quant)
    opt = torch.optim.AdamW(model.parameters(), lr=cfg["learning_rate"])
    step_t = torch.zeros(1, dtype=torch.int64)       # outer iterations done by the RUN (travels with the shared state)
    state = pccl.SharedState([pccl.TensorInfo.from_torch(outer, "outer_params"), pccl.TensorInfo.from_torch(algo.buf, "outer_momentum"),
                              pccl.TensorInfo.from_torch(step_t, "outer_step")])
    ckpt_path = Path(cfg["checkpoint_dir"]) / "ckpt.pt" if cfg["checkpoint_dir"] else None

    if args.resume:
        assert ckpt_path is not None and ckpt_path.exists(), "--resume needs checkpoint_dir/ckpt.pt"
        ckpt = torch.load(ckpt_path, map_location=dev)
        outer.copy_(ckpt["outer_params"])
        algo.buf.copy_(ckpt["outer_momentum"])
        inner_flat.copy_(outer)
        opt.load_state_dict(ckpt["inner_optimizer"])
        step_t.fill_(ckpt["outer_step"])
        algo.steps = int(ckpt["outer_step"])         # the momentum buffer holds something: later steps are not "first" steps
        if hasattr(algo, "applied"):
            algo.applied = algo.steps
        state.revision = int(ckpt["revision"])       # the first synchronisation of a group may carry any revision
        print(f"resumed from {ckpt_path} at outer step {int(step_t)} (shared-state revision {state.revision})", flush=True)

    def save_checkpoint():
        if ckpt_path is None:
            return
        if args.mode == "async":
            algo.flush()
        ckpt_path.parent.mkdir(parents=True, exist_ok=True)
        tmp = ckpt_path.with_suffix(".tmp")
        torch.save({"outer_params": outer, "outer_momentum": algo.buf, "inner_optimizer": opt.state_dict(), "outer_step": int(step_t),
                    "revision": state.revision, "config": cfg}, tmp)
        tmp.replace(ckpt_path)

    prof = Profiler(sync_cuda=dev.type == "cuda")
    gen = torch.Generator().manual_seed(1234 + int(time.time() * 1000) % 1000)
    eval_gen = torch.Generator().manual_seed(99)

    def sync(strategy=pccl.SharedStateSyncStrategy.ENFORCE_POPULAR):
        before = int(step_t)
        info = comm.sync_shared_state(state, strategy)
        if info.rx_bytes:
            inner_flat.copy_(outer)                  # joined late: start from the group's parameters
            if args.mode == "async" and int(step_t) > 0 and before == 0:
                algo.mark_joined_running_group()
        state.push_revision()
        return info

    it, counted, crowd_seen = -1, 0, args.count_from_world <= 0
    while counted < cfg["outer_steps"]:
        it += 1
        with prof.section("membership"):
            if args.mode == "async":
                # shared state is only synchronised when the topology changed (reference: async_diloco.py:419-431,448-451)
                algo.admit_peers(sync)
            else:
                # not in the first iteration: a peer that has just been admitted went through the topology round already, and
                # the others are waiting for it in the shared-state synchronisation (reference: sync_diloco.py, local_iter_num > 1)
                if it > 0 and (comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < 2 or comm.are_peers_pending()):
                    comm.update_topology()
                if comm.get_attribute(pccl.Attribute.PEER_GROUP_WORLD_SIZE) > 1:
                    with prof.section("shared_state"):
                        sync()
        if cfg["eval_interval"] and int(step_t) % cfg["eval_interval"] == 0:
            with prof.section("eval"):
                val = estimate_loss(model, cfg, model_cfg, dev, eval_gen)
                print(f"outer step {int(step_t)}: val loss {val:.4f}", flush=True)
                save_checkpoint()
        with prof.section("inner"):
            for _ in range(cfg["inner_steps"]):
                x = torch.randint(0, model_cfg.vocab_size, (cfg["batch_size"], model_cfg.block_size), generator=gen).to(dev)
                _, loss = model(x, x)
                opt.zero_grad(set_to_none=True)
                loss.backward()
                opt.step()
        with prof.section("outer"):
            if args.mode == "async":
                algo.outer_step(inner_flat, sync, run_step=int(step_t))
            else:
                algo.outer_step(inner_flat)
            step_t += 1
        world = comm.get_attribute(pccl.Attribute.PEER_GROUP_WORLD_SIZE)
        crowd_seen = crowd_seen or world >= args.count_from_world
        counted += 1 if crowd_seen else 0
        plane = comm.ext_last_reduce_stats(0)["plane"] if world > 1 and args.mode == "sync" else "-"
        print(f"outer step {it}: loss {loss.item():.4f} world {world} plane {plane} run step {int(step_t)}", flush=True)
    if args.mode == "async":
        algo.flush()
    save_checkpoint()
    print(prof.report())
    if args.timeline:  # reference: profiler.export_timeline(), sync_diloco.py:538
        prof.export_timeline(args.timeline)
        prof.export_chrome_trace(str(Path(args.timeline).with_suffix(".trace.json")))
    comm.destroy()
    if master:
        master.interrupt()


if __name__ == "__main__":
    main()
