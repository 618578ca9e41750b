"""Tensor-parallel parity on real GPUs (SURVEY.md section 8e "Parity under TP").

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/tp_parity.py [model] [format] [steps]

Every rank builds its shard of the random-init model (piece `rank` of N), rank 0 additionally builds the
SAME global model in one process (`shards=N` concatenates the N pieces) -- then both decode greedily from
the same prompt.  The all-reduce changes the summation order, so the check is the north_star one: identical
greedy tokens for `steps` steps, unless the single-GPU top-2 logit margin at the first divergence is tiny.
Prints one JSON line on rank 0 and exits non-zero on a real mismatch."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch.distributed as dist  # noqa: E402

from pygpukit_b200 import llm  # noqa: E402
from pygpukit_b200._lib import api  # noqa: E402
from pygpukit_b200.llm.tp import torch_allgather, torch_exchange  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "tiny-qwen3"
fmt = sys.argv[2] if len(sys.argv) > 2 else "fp8"
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 64
layers = int(sys.argv[4]) if len(sys.argv) > 4 else None
rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
api.device_set(int(os.environ.get("LOCAL_RANK", "0")))
dist.init_process_group("gloo")          # host plumbing only: the data path is NCCL through libpgk_b200.so
comm = llm.TensorParallel(rank, world).init(exchange=torch_exchange(), allgather=torch_allgather())
max_seq = 256
prompt = [3, 141, 59, 26, 535, 89, 79, 323]


def run(model):
    model.init_decode(max_seq)
    return model.generate(prompt, steps, temperature=0.0)[len(prompt):]


tp_tokens = run(llm.build_random_model(name, fmt, max_seq=max_seq, comm=comm, std=0.05, num_layers=layers))
api.device_sync()
dist.barrier()
ok, detail = True, {}
if rank == 0:
    single = llm.build_random_model(name, fmt, max_seq=max_seq, shards=world, std=0.05, num_layers=layers)
    ref_tokens = run(single)
    same = int(np.argmin([a == b for a, b in zip(tp_tokens, ref_tokens)] + [False]))
    detail = {"model": name, "format": fmt, "world": world, "steps": steps, "identical_prefix": same,
              "tp_tokens": tp_tokens[:16], "single_tokens": ref_tokens[:16]}
    if same < steps:
        # a divergence is only acceptable at a near-tie of the single-GPU logits
        single.init_decode(max_seq)
        ctx = prompt + ref_tokens[:same]
        hidden = single.prefill(ctx) if hasattr(single, "prefill") else None
        logits = single.get_logits(hidden).to_numpy().reshape(-1, single._head.shape[0])[-1] if hidden is not None else None
        detail["tp_token"], detail["single_token"] = tp_tokens[same], ref_tokens[same]
        if logits is not None:
            from pygpukit_b200.core.dtypes import bf16_bits_to_f32
            lf = bf16_bits_to_f32(logits) if logits.dtype == np.uint16 else logits.astype(np.float32)
            top2 = np.sort(lf)[-2:]
            detail["margin_at_divergence"] = float(top2[1] - top2[0])
            ok = detail["margin_at_divergence"] <= 2.0 ** -5
        else:
            ok = False
    detail["allreduce"] = "p2p-nvlink" if comm._p2p else "nccl"
    detail["ok"] = ok
    print(json.dumps(detail), flush=True)
flag = [ok]
dist.broadcast_object_list(flag, src=0)
comm.destroy()
dist.destroy_process_group()
sys.exit(0 if flag[0] else 1)
