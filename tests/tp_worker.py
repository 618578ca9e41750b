"""Tensor-parallel parity worker (SURVEY.md section 8e "Parity under TP"); launched by tests/test_gpu_tp.py as

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tests/tp_worker.py name:format:steps[:layers] [more cases ...]

Every rank builds its shard of the random-init model (piece `rank` of N); rank 0 additionally builds the SAME
global model in one process (`shards=N` concatenates the N pieces).  Both decode greedily from the same prompt.
The all-reduce changes the summation order of the row-parallel projections, so the check is north_star's:
identical greedy tokens for `steps` steps, unless the single-GPU top-2 logit margin at the first divergence is
below 2^-5 (a tie at BF16 logit resolution).  Rank 0 prints one JSON line per case; exit code 1 on a real mismatch."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch.distributed as dist  # noqa: E402

from pygpukit_b200 import llm  # noqa: E402
from pygpukit_b200._lib import api  # noqa: E402
from pygpukit_b200.llm.tp import torch_allgather, torch_exchange  # noqa: E402

rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
api.device_set(int(os.environ.get("LOCAL_RANK", "0")))
api.set_pdl(int(os.environ.get("PGK_PDL", "1")))
dist.init_process_group("gloo")          # host plumbing only: the data path is NCCL / NVLink through libpgk_b200.so
comm = llm.TensorParallel(rank, world).init(exchange=torch_exchange(), allgather=torch_allgather())
MAX_SEQ = 256
PROMPT = [3, 141, 59, 26, 535, 89, 79, 323]


def run(model, steps):
    model.init_decode(MAX_SEQ)
    return model.generate(PROMPT, steps, temperature=0.0)[len(PROMPT):]


def step_state(model):
    """'step' when the decode ran as the one-launch step kernel (in-kernel NVLink exchange), else why not."""
    prog = model.step_program()
    if prog is None:
        return "per-op: " + str(model.step_unsupported_reason)
    return "step" if prog.error() == 0 else f"step ERROR {prog.error()}"


def moe_layer_case(case, fmt):
    """Expert parallelism in isolation: the SAME input rows through the sharded MoE layer (experts split over the ranks,
    un-combined down-projection rows exchanged) and through the one-process layer must give the SAME BITS -- one
    non-zero contributor per row makes the exchange exact, the weighted combine then runs identically everywhere."""
    from pygpukit_b200.core.factory import from_numpy
    tp_model = llm.build_random_model("tiny-moe", fmt, max_seq=MAX_SEQ, comm=comm, std=0.05)
    rng = np.random.default_rng(11)
    H = tp_model.config.hidden_size
    outs = []
    for T in (1, 5):
        x = ((rng.standard_normal((T, H)) * 0.5).astype(np.float32).view(np.uint32) >> 16).astype(np.uint16)
        res = ((rng.standard_normal((T, H)) * 0.5).astype(np.float32).view(np.uint32) >> 16).astype(np.uint16)
        y = tp_model.blocks[1].mlp.forward(from_numpy(x), residual=from_numpy(res), comm=comm)
        outs.append((x, res, y.to_numpy()))
    api.device_sync()
    dist.barrier()
    ok, detail = True, {}
    if rank == 0:
        single = llm.build_random_model("tiny-moe", fmt, max_seq=MAX_SEQ, shards=world, std=0.05)
        same = []
        for x, res, y in outs:
            y1 = single.blocks[1].mlp.forward(from_numpy(x), residual=from_numpy(res)).to_numpy()
            same.append(bool(np.array_equal(y, y1)))
        ok = all(same)
        detail = {"case": case, "world": world, "format": fmt, "bit_identical_T1_T5": same, "ok": ok, "steps": 0,
                  "identical_prefix": 0}
        print(json.dumps(detail), flush=True)
    return ok


def linear_allreduce_case(case, fmt):
    """Row-parallel GEMV + all-reduce + residual in ONE launch (pgk_gemv_*_allreduce_bf16) against the two launches it
    replaces (GEMV -> pgk_p2p_ll_allreduce_add_bf16): the SAME BITS on every rank, for several shapes, repeated so that both
    parity halves of the LL regions and a changing hidden state are exercised."""
    from pygpukit_b200.core.factory import from_numpy
    from pygpukit_b200.llm.random_init import normal_bf16, quantize_fp8, quantize_nvf4
    rng = np.random.default_rng(100 + rank)
    same = []
    for K, N in ((1024, 512), (2048, 4096), (3584, 8192), (512, 1536)):
        if fmt == "fp8":
            lin = quantize_fp8(normal_bf16((N, K), 40 + rank, 0.05))
        else:
            lin = quantize_nvf4(normal_bf16((K, N), 40 + rank, 0.05))
        assert comm.fuses_allreduce(lin, 1), (fmt, K, N)
        hid0 = ((np.random.default_rng(7).standard_normal((1, N)) * 0.5).astype(np.float32).view(np.uint32) >> 16).astype(np.uint16)
        h_two, h_one = from_numpy(hid0.copy()), from_numpy(hid0.copy())
        part = from_numpy(np.zeros((1, N), np.uint16))
        for it in range(5):
            x = from_numpy(((rng.standard_normal((1, K))).astype(np.float32).view(np.uint32) >> 16).astype(np.uint16))
            lin.forward(x, out=part)
            comm.allreduce_add(part, h_two)
            lin.forward_allreduce(x, h_one)
            api.device_sync()
            same.append(bool(np.array_equal(h_two.to_numpy(), h_one.to_numpy())))
    flags = [None] * world
    dist.all_gather_object(flags, all(same))
    ok = all(flags)
    if rank == 0:
        print(json.dumps({"case": case, "world": world, "format": fmt, "bit_identical": same, "all_ranks": flags, "ok": ok,
                          "steps": 0, "identical_prefix": 0}), flush=True)
    return ok


def router_tie_before(single, tokens, upto):
    """True when, while the one-process model decodes `tokens[:upto + 1]` after the prompt, some MoE layer's router sees
    its k-th and (k+1)-th logits equal or ADJACENT in BF16: the tensor-parallel run (different summation order in the
    attention all-reduce) may then route that token to a different expert -- a tie, but a discontinuous one."""
    from pygpukit_b200.core.dtypes import bf16_bits_to_f32
    single.use_step_program = False
    single.init_decode(MAX_SEQ)
    seq = PROMPT + tokens[:upto + 1]
    k = single.config.num_experts_per_tok
    for pos, t in enumerate(seq):
        single._decode_step_fixed_cache(t, pos, pos + 1)
        for blk in single.blocks:
            bufs = getattr(blk.mlp, "_bufs", {}).get(1)
            if bufs is None:
                continue
            lg = np.sort(bf16_bits_to_f32(bufs["logits"].to_numpy().reshape(-1)))[::-1]
            gap, top = float(lg[k - 1] - lg[k]), max(abs(float(lg[k - 1])), 1e-30)
            if gap <= 2.0 ** (np.floor(np.log2(top)) - 7):
                return True
    return False


all_ok = True
for case in sys.argv[1:]:
    f = case.split(":")
    if f[0] in ("moe-layer", "linear-allreduce"):
        ok = moe_layer_case(case, f[1]) if f[0] == "moe-layer" else linear_allreduce_case(case, f[1])
        flag = [ok]
        dist.broadcast_object_list(flag, src=0)
        all_ok = all_ok and flag[0]
        continue
    name, fmt, steps = f[0], f[1], int(f[2])
    layers = int(f[3]) if len(f) > 3 and f[3] else None
    tp_model = llm.build_random_model(name, fmt, max_seq=MAX_SEQ, comm=comm, std=0.05, num_layers=layers)
    if len(f) > 4:                          # fifth field: "step" / "perop" pins the decode path
        tp_model.use_step_program = f[4] == "step"
    tp_tokens = run(tp_model, steps)
    path = step_state(tp_model)
    api.device_sync()
    dist.barrier()
    ok, detail = True, {}
    if rank == 0:
        single = llm.build_random_model(name, fmt, max_seq=MAX_SEQ, shards=world, std=0.05, num_layers=layers)
        ref_tokens = run(single, steps)
        same = int(np.argmin([a == b for a, b in zip(tp_tokens, ref_tokens)] + [False]))
        detail = {"case": case, "model": name, "format": fmt, "world": world, "steps": steps, "identical_prefix": same,
                  "tp_tokens": tp_tokens[:16], "single_tokens": ref_tokens[:16]}
        if same < steps:        # a divergence is only acceptable at a near-tie of the single-GPU logits
            single.init_decode(MAX_SEQ)
            hidden = single.prefill(PROMPT + ref_tokens[:same])
            logits = single.get_logits(hidden).to_numpy().reshape(-1, single._head.shape[0])[-1]
            from pygpukit_b200.core.dtypes import bf16_bits_to_f32
            top2 = np.sort(bf16_bits_to_f32(logits))[-2:]
            detail.update(tp_token=tp_tokens[same], single_token=ref_tokens[same], margin_at_divergence=float(top2[1] - top2[0]))
            ok = detail["margin_at_divergence"] <= 2.0 ** -5
            if not ok and single.config.is_moe:
                # dense parts differ by summation order only, but a router near-tie turns that into another expert
                detail["router_tie_before_divergence"] = router_tie_before(single, ref_tokens, same)
                ok = detail["router_tie_before_divergence"]
        detail["allreduce"] = "p2p-nvlink" if comm._p2p else "nccl"
        detail["decode_path"] = path
        if "ERROR" in path or (len(f) > 4 and f[4] == "step" and path != "step"):
            ok = False
        detail["ok"] = ok
        print(json.dumps(detail), flush=True)
        del single
    flag = [ok]
    dist.broadcast_object_list(flag, src=0)
    all_ok = all_ok and flag[0]
comm.destroy()
dist.destroy_process_group()
sys.exit(0 if all_ok else 1)
