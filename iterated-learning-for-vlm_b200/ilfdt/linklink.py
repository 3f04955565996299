"""Mirror of prototype/linklink/__init__.py: thin aliases over torch.distributed (NCCL over NVLink on
the GPU box, gloo in CPU tests). The hot-path collectives themselves are issued by
ilfdt.functional.AllGather (all_gather_into_tensor / reduce_scatter_tensor)."""
import functools
import os

import torch
import torch.distributed as dist

allreduce = dist.all_reduce
allgather = dist.all_gather
broadcast = dist.broadcast
synchronize = torch.cuda.synchronize
init_process_group = dist.init_process_group
allreduce_async = functools.partial(dist.all_reduce, async_op=True)


def get_rank():
    return int(os.environ.get('RANK', 0))


def get_world_size():
    return int(os.environ.get('WORLD_SIZE', 1))


def get_local_rank():
    return int(os.environ.get('LOCAL_RANK', 0))


def barrier():
    """linklink/__init__.py:30-34 all-reduces an IntTensor and reads it back on the host (a full host
    sync three times per step). dist.barrier() has the same ordering effect without the read-back."""
    if get_world_size() > 1 and dist.is_initialized():
        dist.barrier()
