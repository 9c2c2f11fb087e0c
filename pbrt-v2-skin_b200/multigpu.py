"""Host-side plumbing of the multi-GPU path when every GPU has its own process (torchrun): which points a rank computes, and
how rank 0's NCCL id reaches the others.  The exchanges themselves -- all-gather of the irradiance points, sum of the films --
are NCCL calls inside libskincuda.so (csrc/skc_comm.cu, skc.SkinComm); nothing here touches the data.  The same partition
is used by the one-process form (`pbrt --gpus N`, host/skin_render.cpp RenderSceneRank)."""
import numpy as np


def shard_range(n, rank, world):
    """Contiguous shard [first, first+count) of n items and the padded shard size: every shard but the last has
    ceil(n / world) items (host/skin_render.cpp RenderSceneRank uses the same rule)."""
    per = (n + world - 1) // world
    first = min(rank * per, n)
    return first, min(per, n - first), per


def exchange_unique_id(dist, rank, make_id):
    """rank 0 creates the communicator id (make_id() -> 128 bytes); every rank returns the same bytes.
    dist: an initialised torch.distributed process group on a CPU backend (gloo): the launcher's rendezvous, used for these
    128 bytes and nothing else."""
    import torch
    t = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        t = torch.from_numpy(np.frombuffer(make_id(), np.uint8).copy())
    dist.broadcast(t, src=0)
    return t.numpy().tobytes()
