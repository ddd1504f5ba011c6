"""torchrun worker of tests/test_ddp_gpu.py: one rank per GPU, NCCL.  Prints one JSON line per bucket-group size on
rank 0 (plaincv_b200.selfcheck.ddp_gradient_check: N-rank gradients vs the whole global batch on one rank)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    import faulthandler
    if int(os.environ.get("PCV_TEST_DUMP_AFTER", "0")) > 0:  # debugging aid: where is every thread after N seconds?
        faulthandler.dump_traceback_later(int(os.environ["PCV_TEST_DUMP_AFTER"]), exit=True)
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from plaincv_b200.batch import synthetic_batch
    from plaincv_b200.model import ModelConfig, Transformer
    from plaincv_b200.selfcheck import ddp_gradient_check
    use_doc = os.environ.get("PCV_TEST_DOC", "0") == "1"
    n_layers = int(os.environ.get("PCV_TEST_LAYERS", "9"))
    T, B, V = 1024, 2, 50280
    cfg = ModelConfig(vocab_size=V, seq_len=T, dim=1024, expand=8 / 3, n_layers=n_layers, n_heads=16)
    model = Transformer(cfg, device=dev, seed=0)  # same seed on every rank: identical replicas
    batches = [synthetic_batch(B, T, V, seed=500 + k, doc_mean=300 if use_doc else None) for k in range(world)]
    for lpb in (1, 4, 8):
        print(f"rank {rank}: layers_per_bucket {lpb}", file=sys.stderr, flush=True)
        res = ddp_gradient_check(model, T, use_doc, batches, rank, world, layers_per_bucket=lpb)
        res["layers_per_bucket"] = lpb
        if rank == 0:
            print("DDPCHECK " + json.dumps(res), flush=True)
    # No destroy_process_group(): tearing the NCCL communicator down while CUDA graphs that captured its collectives
    # are still alive blocks forever (seen on 2 x B200, torch 2.11 / NCCL 2.28).  Everything is flushed; leave.
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
