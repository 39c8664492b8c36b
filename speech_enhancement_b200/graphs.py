"""CUDA-graph replay of a caller model's inference forward.

The chained forward of DCCRN is ~60 launches of this package plus the PyTorch glue around the LSTM;
at small per-GPU batches (BASELINE configs[0] at batch 1, configs[2] at 16 per GPU) the step is bound by
launch latency, not by the kernels.  Every launch of this package goes on torch's current stream with
no host synchronisation, tensor maps are encoded on the host before the launch and weight operands are
cached per parameter version, so the whole forward can be captured once per input shape and replayed.
"""
import torch


class GraphedInference:
    """`GraphedInference(model)(x)` == `model(x)` under eval / no_grad, replayed from a CUDA graph.

    One graph per input shape (captured on first use after `warmup` eager calls, which also build the
    cached weight operands).  The outputs are static buffers owned by the graph: they are overwritten by
    the next call with the same shape (clone them to keep them).  If any parameter or buffer of the model
    changes (optimizer step, load_state_dict), call `reset()`.
    """

    def __init__(self, model, warmup=2):
        self.model = model.eval()
        self.warmup = warmup
        self._graphs = {}

    def reset(self):
        self._graphs.clear()

    def _capture(self, x):
        static_in = x.clone()
        side = torch.cuda.Stream(device=x.device)
        side.wait_stream(torch.cuda.current_stream(x.device))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(self.warmup):
                self.model(static_in)
        torch.cuda.current_stream(x.device).wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(graph):
            static_out = self.model(static_in)
        return graph, static_in, static_out

    def __call__(self, x):
        if not x.is_cuda:
            raise RuntimeError('GraphedInference: CUDA input required')
        key = (tuple(x.shape), x.dtype, x.device)
        entry = self._graphs.get(key)
        if entry is None:
            entry = self._graphs[key] = self._capture(x)
        graph, static_in, static_out = entry
        static_in.copy_(x, non_blocking=True)
        graph.replay()
        return static_out
