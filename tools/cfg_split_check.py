#!/usr/bin/env python
"""CFG-branch split (three GPUs, one all-gather per guided step) against the single-GPU sampler: same result, lower
latency.  Launch with >= 3 ranks:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 tools/cfg_split_check.py [--tiny]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tiny", action="store_true")
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    group = dist.new_group([0, 1, 2])
    from echo_tts_b200 import schema as S, serving
    from echo_tts_b200.inference import load_model_from_state_dict
    from echo_tts_b200.sampler import sample_euler_cfg_independent_guidances as sample
    import bench
    cfg = S.ECHO_TINY if args.tiny else S.ECHO_BASE
    model = load_model_from_state_dict(S.synth_state_dict(S.dit_schema(cfg, blockwise=False), init="torch"), cfg=cfg)
    ids, mask, spk, smask = [t.cuda() for t in bench.synth_inputs(1)]
    kw = dict(bench.KW)
    if args.tiny:
        kw.update(sequence_length=96, num_steps=8)
    noise = torch.randn(1, kw["sequence_length"], 80, generator=torch.Generator().manual_seed(3))
    out = {}
    if rank < 3:
        def timed(fn):
            fn()
            torch.cuda.synchronize(); dist.barrier(group)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.reps):
                r = fn()
            e1.record()
            torch.cuda.synchronize(); dist.barrier(group)
            return r, serving.max_over_ranks_ms(e0.elapsed_time(e1) / args.reps, torch.device("cuda", local), group)
        split, ms_split = timed(lambda: serving.sample_euler_cfg_split(model, spk, smask, ids, mask, 0, group=group,
                                                                        _noise=noise, **kw))
        split_p2p, ms_p2p = timed(lambda: serving.sample_euler_cfg_split(model, spk, smask, ids, mask, 0, group=group,
                                                                          exchange="p2p", _noise=noise, **kw))
        os.environ["ECHO_CUDA_GRAPH"] = "0"
        single_eager, ms_eager = timed(lambda: sample(model, spk, smask, ids, mask, 0, _noise=noise, **kw))
        os.environ["ECHO_CUDA_GRAPH"] = "1"
        single, ms_graph = timed(lambda: sample(model, spk, smask, ids, mask, 0, _noise=noise, **kw))
        gathered = [torch.empty_like(split) for _ in range(3)]
        dist.all_gather(gathered, split.contiguous(), group=group)
        if rank == 0:
            rel = float((split - single_eager).norm() / single_eager.norm())
            same = all(torch.equal(gathered[0], g) for g in gathered[1:])
            out = dict(check="cfg_split", arch="tiny" if args.tiny else "base", steps=kw["num_steps"], S=kw["sequence_length"],
                       rel_l2_vs_single_gpu=rel, ranks_bit_identical=same, ms_split_3gpu=ms_split,
                       p2p_equals_nccl=bool(torch.equal(split_p2p, split)), ms_split_3gpu_p2p=ms_p2p,
                       ms_single_gpu_eager=ms_eager, ms_single_gpu_graph=ms_graph)
            assert rel < 1e-2 and same and out["p2p_equals_nccl"], out      # bf16 paths with different GEMM summation splits
    dist.barrier()
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
