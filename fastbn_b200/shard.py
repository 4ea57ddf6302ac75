"""Multi-GPU driver of the junction-tree path: one process per GPU (torchrun), test cases cut into contiguous slices
[r*n/G, (r+1)*n/G), the plan replicated on every GPU, NO data-path collective; the only exchange is the final
all-reduce of {n_correct, n_total, sum_mse, sum_hd} — NCCL on GPUs, gloo in the CPU tests.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \\
      -m fastbn_b200.shard --xml bn.xml --cases cases.libsvm --pt golden_pt

prints the reference's result lines (average MSE / average HD / accuracy) on rank 0."""
import argparse
import os

import numpy as np


def case_slice(n_cases, rank, world):
    """Contiguous slice of rank `rank` (the same split the C++ mirror uses per device, host/junction_tree.cpp)."""
    return (n_cases * rank) // world, (n_cases * (rank + 1)) // world


def slice_csr(ev_off, ev_vars, ev_vals, c0, c1):
    """CSR evidence of cases [c0, c1) with offsets rebased to 0."""
    e0, e1 = int(ev_off[c0]), int(ev_off[c1])
    return ev_off[c0:c1 + 1] - e0, ev_vars[e0:e1], ev_vals[e0:e1]


def comm_from_process_group(device):
    """fbn_comm over the ranks of the default torch.distributed group: rank 0 makes the NCCL unique id (fbn_comm_unique_id),
    the process group's store carries its 128 bytes to the other ranks, every rank calls fbn_comm_create."""
    import torch.distributed as dist
    from . import capi
    rank, world = dist.get_rank(), dist.get_world_size()
    box = [capi.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return capi.Comm(device, rank, world, box[0])


def reduce_scores(out4, device=None, comm=None):
    """Sum of the per-rank {n_correct, n_total, sum_mse, sum_hd}: through fbn_reduce (one ncclAllReduce behind the C ABI) when
    a communicator is given, else over the default torch.distributed group (gloo in the CPU tests; no-op without a group)."""
    if comm is not None:
        v = comm.reduce([out4["n_correct"], out4["n_total"], out4["sum_mse"], out4["sum_hd"]])
        return dict(n_correct=int(round(v[0])), n_total=int(round(v[1])), sum_mse=float(v[2]), sum_hd=float(v[3]))
    import torch
    import torch.distributed as dist
    t = torch.tensor([out4["n_correct"], out4["n_total"], out4["sum_mse"], out4["sum_hd"]], dtype=torch.float64, device=device or "cpu")
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    v = t.cpu().numpy()
    return dict(n_correct=int(round(v[0])), n_total=int(round(v[1])), sum_mse=float(v[2]), sum_hd=float(v[3]))


def evaluate_sharded(tree, ev_off, ev_vars, ev_vals, golden, truths, rank, world, device, comm=None):
    """This rank's slice through fbn_evaluate on cuda:`device`, then the reduction of the four scores (fbn_reduce when `comm`
    is an fbn_comm, else torch.distributed). Returns (labels of the slice, scores summed over the ranks)."""
    import torch
    from . import capi
    c0, c1 = case_slice(len(ev_off) - 1, rank, world)
    off, ev, ew = slice_csr(ev_off, ev_vars, ev_vals, c0, c1)
    plan = capi.Plan(tree, device=device)
    labels, _, sc, _ = plan.evaluate(off, ev, ew, golden[c0:c1], truths[c0:c1])
    plan.close()
    return labels, reduce_scores(sc, torch.device("cuda", device), comm=comm)


def main():
    import torch
    import torch.distributed as dist
    from . import hostapi
    ap = argparse.ArgumentParser()
    ap.add_argument("--xml", required=True)
    ap.add_argument("--cases", required=True)
    ap.add_argument("--pt", required=True)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    m = hostapi.HostModel(a.xml, a.cases)
    tree = m.tree()
    off, ev, ew, truths = m.cases()
    golden = m.golden(a.pt)
    comm = comm_from_process_group(local) if world > 1 else None
    labels, sc = evaluate_sharded(tree, off, ev, ew, golden, truths, rank, world, local, comm=comm)
    if comm is not None:
        comm.close()
    if rank == 0:
        print("average MSE = %g" % (sc["sum_mse"] / sc["n_total"]))
        print("average HD = %g" % (sc["sum_hd"] / sc["n_total"]))
        print("accuracy = %g" % (sc["n_correct"] / sc["n_total"]))
    m.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
