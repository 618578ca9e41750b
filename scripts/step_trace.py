"""Per-op timing INSIDE the one-launch step kernel (pgk_step_trace: globaltimer at which CTA 0 starts each op).

    python scripts/step_trace.py --model llama3-8b --format fp8 --context 4096 [--steps 16]

Prints, per op slot of a layer (qkv, attention, o_proj, gate/up, down) and for the lm_head, the mean duration over layers
and steps with the bytes the op streams and the resulting GB/s.  A duration includes the grid barrier that ends the op."""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from pygpukit_b200 import llm  # noqa: E402
from pygpukit_b200._lib import api  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3-8b")
    ap.add_argument("--format", default="fp8")
    ap.add_argument("--context", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--layers", type=int, default=None)
    a = ap.parse_args()
    api.set_pdl(1)
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    comm = None
    if world > 1:                     # torchrun: tensor-parallel shards, the exchanges happen inside the step kernel
        import torch.distributed as dist
        from pygpukit_b200.llm.tp import torch_allgather, torch_exchange
        api.device_set(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group("gloo")
        comm = llm.TensorParallel(rank, world).init(exchange=torch_exchange(), allgather=torch_allgather())
    m = llm.build_random_model(a.model, a.format, max_seq=a.context, num_layers=a.layers, comm=comm)
    m.use_step_program = True
    m.init_decode(a.context)
    for blk in m.blocks:
        for c in (blk.attn._k_cache, blk.attn._v_cache):
            api.fill_normal_bf16(c.device_ptr, c.size, 7, 0.1, None)
    m.init_decode_graph()
    prog = m.step_program()
    if prog is None:
        raise SystemExit(f"no step program: {m.step_unsupported_reason}")
    if rank != 0:
        sys.stdout = open(os.devnull, "w")
    print("info", prog.info(), "world", world)
    L = len(m.blocks)
    p0 = a.context - a.steps - 4
    m.set_decode_state(123, p0)
    acc = []
    for _ in range(a.steps):
        m.decode_graph_replay()
        acc.append(prog.trace().astype(np.int64))
    assert prog.error() == 0
    tr = np.array(acc[2:], np.float64) / 1e3             # us, [steps, n_ops + 1, 8]
    d = np.diff(tr[:, :, 0], axis=1)                     # op start -> next op start
    staged = tr[:, :-1, 1] - tr[:, :-1, 0]               # barrier passed -> activations staged / attention prologue done
    w0 = tr[:, :-1, 2] - tr[:, :-1, 1]                   # -> warp 0 finished its items
    cta0 = tr[:, :-1, 4] - tr[:, :-1, 0]                 # -> CTA 0 arrived at the closing barrier
    wait = tr[:, 1:, 0] - tr[:, :-1, 4]                  # closing barrier: CTA 0's arrival -> released
    per_layer = d[:, :5 * L].reshape(-1, L, 5)
    brk = [x[:, :5 * L].reshape(-1, L, 5).mean(axis=(0, 1)) for x in (staged, w0, cta0, wait)]
    blk = m.blocks[0]
    cfg = m.config
    H = cfg.hidden_size
    kvb = 2 * blk.attn.num_kv_heads * a.context * 128 * 2
    names = ["qkv (+norm)", "attention (+fold)", "o_proj (+res)", "gate/up (+norm, swiglu)", "down (+res)"]
    byts = [blk.attn.qkv_proj.weight_bytes, kvb, blk.attn.o_proj.weight_bytes,
            blk.mlp.weight_bytes - blk.mlp.down_proj.weight_bytes, blk.mlp.down_proj.weight_bytes]
    tot_us, tot_b = 0.0, 0.0
    for i, (n, b) in enumerate(zip(names, byts)):
        us = per_layer[:, :, i].mean()
        tot_us += us
        tot_b += b
        print(f"{n:26s} {b / 1e6:8.2f} MB {us:7.2f} us {b / us / 1e3:7.0f} GB/s   staged {brk[0][i]:5.2f}  warp0 items {brk[1][i]:5.2f}  "
              f"CTA0 busy {brk[2][i]:5.2f}  barrier wait {brk[3][i]:5.2f}")
    print(f"{'layer':26s} {tot_b / 1e6:8.2f} MB {tot_us:7.2f} us {tot_b / tot_us / 1e3:7.0f} GB/s")
    hb = m._head.nbytes
    us = d[:, 5 * L].mean()
    print(f"{'lm_head (+norm, argmax)':26s} {hb / 1e6:8.2f} MB {us:7.2f} us {hb / us / 1e3:7.0f} GB/s")
    print(f"step {d.sum(axis=1).mean():.1f} us (CTA 0, first op start .. end)")
    for nm, idx in (("qkv", 0), ("gate/up", 3), ("o_proj", 2), ("down", 4)):
        o = tr[:, idx:5 * L:5, :]
        print(f"staging of {nm}: loads+raw {(o[:, :, 5] - o[:, :, 0]).mean():.2f}  sumsq {(o[:, :, 6] - o[:, :, 5]).mean():.2f}  "
              f"norm+max {(o[:, :, 7] - o[:, :, 6]).mean():.2f}  planes {(o[:, :, 1] - o[:, :, 7]).mean():.2f}   (no norm: only [7]-[0] = "
              f"{(o[:, :, 7] - o[:, :, 0]).mean():.2f} and planes are meaningful)")
    at = tr[:, 1:5 * L:5, :]                              # attention ops
    print("attention: prologue %.2f  units(warp0) %.2f  fold %.2f  inner barrier %.2f  combine+closing %.2f" % (
        (at[:, :, 1] - at[:, :, 0]).mean(), (at[:, :, 2] - at[:, :, 1]).mean(), (at[:, :, 3] - at[:, :, 2]).mean(),
        (at[:, :, 6] - at[:, :, 3]).mean(), (tr[:, 2:5 * L + 1:5, 0] - at[:, :, 6]).mean()))
    _ = H


if __name__ == "__main__":
    main()
