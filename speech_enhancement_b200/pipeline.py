"""Host-to-host enhancement with the copies overlapped with the kernels.

`model(x)` on a CUDA tensor is the drop-in call; a serving loop, though, starts from pinned host buffers and
wants the enhanced waveform back on the host.  `HostPipeline` runs that loop on three streams: the H2D copy of
batch i+1 and the D2H copy of result i-1 travel while batch i is being enhanced (two device input buffers and
two device output buffers per shape, ordered with events; no allocation after the first batch of a shape).
The arithmetic is exactly `model(x)[1]`.
"""
import torch


class HostPipeline:
    """`p = HostPipeline(model); p.submit(host_x, host_out); ...; p.synchronize()`.

    `host_x` [B, L] and `host_out` [B, L'] are pinned CPU tensors; `host_out` holds the enhanced waveforms once
    `synchronize()` (or a later `submit` that reuses the same slot, two submits on) has returned.  Results are
    written in submission order.
    """

    def __init__(self, model, device=None):
        self.model = model.eval()
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        if self.device.type != 'cuda':
            raise RuntimeError('HostPipeline: the model must live on a CUDA device')
        self.h2d = torch.cuda.Stream(device=self.device)
        self.d2h = torch.cuda.Stream(device=self.device)
        self._slots = {}
        self._n = 0

    def _slot(self, shape, i):
        key = (tuple(shape), i)
        s = self._slots.get(key)
        if s is None:
            s = self._slots[key] = {'x': torch.empty(shape, device=self.device, dtype=torch.float32), 'y': None,
                                    'in_ready': torch.cuda.Event(), 'in_free': None,
                                    'out_ready': torch.cuda.Event(), 'out_free': None}
        return s

    def output_length(self, host_x):
        """length of the enhanced waveform for this input shape (runs one batch of the same shape on first use)"""
        with torch.no_grad():
            return self.model(torch.zeros(host_x.shape, device=self.device))[1].shape[1]

    def submit(self, host_x, host_out):
        if not (host_x.is_pinned() and host_out.is_pinned()):
            raise RuntimeError('HostPipeline.submit: host buffers must be pinned (tensor.pin_memory())')
        compute = torch.cuda.current_stream(self.device)
        s = self._slot(host_x.shape, self._n & 1)
        self._n += 1
        with torch.cuda.stream(self.h2d):
            if s['in_free'] is not None:
                self.h2d.wait_event(s['in_free'])            # the forward that last read this buffer is done
            s['x'].copy_(host_x, non_blocking=True)
            s['in_ready'].record(self.h2d)
        compute.wait_event(s['in_ready'])
        if s['out_free'] is not None:
            compute.wait_event(s['out_free'])                # the D2H copy that last read this output is done
        with torch.no_grad():
            wav = self.model(s['x'])[1]
        if s['y'] is None or s['y'].shape != wav.shape:
            s['y'] = torch.empty_like(wav)
        s['y'].copy_(wav)                                    # static output buffer: the copy engine reads it later
        s['in_free'] = torch.cuda.Event()
        s['in_free'].record(compute)
        s['out_ready'].record(compute)
        with torch.cuda.stream(self.d2h):
            self.d2h.wait_event(s['out_ready'])
            host_out.copy_(s['y'], non_blocking=True)
            s['out_free'] = torch.cuda.Event()
            s['out_free'].record(self.d2h)

    def join(self):
        """order the current stream behind every copy issued so far (so that an event recorded next covers them)"""
        compute = torch.cuda.current_stream(self.device)
        compute.wait_stream(self.h2d)
        compute.wait_stream(self.d2h)

    def synchronize(self):
        self.h2d.synchronize()
        torch.cuda.current_stream(self.device).synchronize()
        self.d2h.synchronize()
