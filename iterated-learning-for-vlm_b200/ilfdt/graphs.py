"""CUDA-graph capture of a whole training step of the hot path.

At small per-GPU batches (8 GPUs x 4096 rows) one step is ~200 kernel launches, two autograd graphs and three NCCL
calls issued from Python in ~16 ms: the host, not the GPU, sets the pace. Every launch of this library is
graph-safe by construction -- no allocation, no host synchronisation, device-side extents for the ragged text
path (text_buckets.cu) -- so forward + backward of a step can be captured once and replayed.

    step = GraphedStep(fn, (q_img, q_txt, mask))     # fn(*inputs) -> tensor(s); runs fwd + bwd, reads only `inputs`
    loss = step()                                    # replay on the captured inputs
    loss = step(new_q_img, new_q_txt, new_mask)      # copy into the captured input buffers, then replay

Parameters and their .grad tensors keep their addresses between replays (fn must set grads to None itself, as
optimizers with set_to_none do, so that backward writes fresh pool memory that the replay reuses)."""
import torch


class GraphedStep:
    def __init__(self, fn, inputs, warmup=2, pool=None):
        self.inputs = tuple(inputs)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):          # warm-up off the default stream, as graph capture requires
            for _ in range(warmup):
                fn(*self.inputs)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, pool=pool):
            self.outputs = fn(*self.inputs)

    def pool(self):
        return self.graph.pool()

    def release(self):
        """Destroy the captured graph and drop its tensors. A graph that captured NCCL collectives must be released
        BEFORE dist.destroy_process_group(): tearing the communicator down under a live graph can block."""
        if self.graph is not None:
            torch.cuda.synchronize()
            self.graph.reset()
        self.graph, self.outputs, self.inputs = None, None, ()

    def __call__(self, *new_inputs):
        if new_inputs:
            with torch.no_grad():
                for dst, src in zip(self.inputs, new_inputs):
                    if dst.data_ptr() != src.data_ptr():
                        dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.outputs
