"""Host-to-host enhancement: copies overlapped with the kernels, PCM16 on the wire, long clips in chunks.

`model(x)` on a CUDA tensor is the drop-in call; a serving loop, though, starts from pinned host buffers -- fp32,
or the int16 PCM_S 16 samples of the wav files the reference reads and writes (audio_dataset.py:12-21,
mix_audio.py:144-146, trainer.py:255-262) -- and wants the enhanced waveform back on the host.

`HostPipeline` runs that loop on three streams: the H2D copy of batch i+1 and the D2H copy of result i-1 travel
while batch i is being enhanced (two device input buffers and two device output buffers per shape, ordered with
events; no allocation after the first batch of a shape).  The arithmetic is exactly `model(x)[1]`.  With
`dtype=torch.int16` the buffers on both sides are int16: half the H2D and D2H bytes, the int16 -> fp32 conversion
fused into the STFT's stage load and clamp + round + saturate fused into the iSTFT's overlap-add epilogue.

`stft_in_chunks` / `istft_in_chunks` / `roundtrip_in_chunks` process clips longer than a chunk budget block by block
with `ceil(win / hop) - 1` frames of context, equal to the whole-clip call to fp32 rounding (the same frames enter
the same sums); `enhance_long` applies a whole model to a long clip in overlapping chunks (exact for frame-local models,
an approximation with `context_s` seconds of warm-up for the recurrent ones).
"""
import torch

from . import _abi


class HostPipeline:
    """`p = HostPipeline(model); p.submit(host_x, host_out); ...; p.synchronize()`.

    `host_x` [B, L] and `host_out` [B, L'] are pinned CPU tensors of `dtype` (fp32, or int16 for PCM16 on the wire);
    `host_out` holds the enhanced waveforms once `synchronize()` (or a later `submit` that reuses the same slot, two
    submits on) has returned.  Results are written in submission order.
    """

    def __init__(self, model, device=None, dtype=torch.float32):
        self.model = model.eval()
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        if self.device.type != 'cuda':
            raise RuntimeError('HostPipeline: the model must live on a CUDA device')
        if dtype not in (torch.float32, torch.int16):
            raise RuntimeError('HostPipeline: dtype must be torch.float32 or torch.int16 (PCM_S 16)')
        self.dtype = dtype
        self.h2d = torch.cuda.Stream(device=self.device)
        self.d2h = torch.cuda.Stream(device=self.device)
        self._slots = {}
        self._n = 0

    def _forward(self, x):
        """model(x)[1] in the pipeline's sample type"""
        istft = getattr(self.model, 'istft', None)
        if self.dtype != torch.int16 or istft is None or not hasattr(istft, 'pcm16_out'):
            with torch.no_grad():
                wav = self.model(x)[1]
            if self.dtype == torch.int16 and wav.dtype != torch.int16:
                from .functional import pcm16_from_float
                wav = pcm16_from_float(wav)
            return wav
        # PCM16 out of the iSTFT epilogue.  A forward that is not one of this package's chained ones clamps the
        # waveform itself (the reference's clamp_(-1, 1) would zero int16 samples), so those get fp32 + one
        # conversion pass instead.
        chained = getattr(self.model, '_se_chained', None) is not None or hasattr(self.model, '_forward_fused')
        old = istft.pcm16_out
        istft.pcm16_out = bool(chained)
        try:
            with torch.no_grad():
                wav = self.model(x)[1]
        finally:
            istft.pcm16_out = old
        if wav.dtype != torch.int16:
            from .functional import pcm16_from_float
            wav = pcm16_from_float(wav)
        return wav

    def _slot(self, shape, i):
        key = (tuple(shape), i)
        s = self._slots.get(key)
        if s is None:
            s = self._slots[key] = {'x': torch.empty(shape, device=self.device, dtype=self.dtype), 'y': None,
                                    'in_ready': torch.cuda.Event(), 'in_free': None,
                                    'out_ready': torch.cuda.Event(), 'out_free': None}
        return s

    def output_length(self, host_x):
        """length of the enhanced waveform for this input shape (runs one batch of the same shape on first use)"""
        return self._forward(torch.zeros(host_x.shape, device=self.device, dtype=self.dtype)).shape[1]

    def submit(self, host_x, host_out):
        if not (host_x.is_pinned() and host_out.is_pinned()):
            raise RuntimeError('HostPipeline.submit: host buffers must be pinned (tensor.pin_memory())')
        if host_x.dtype != self.dtype or host_out.dtype != self.dtype:
            raise RuntimeError(f'HostPipeline.submit: buffers must be {self.dtype} (got {host_x.dtype} / {host_out.dtype})')
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
        wav = self._forward(s['x'])
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


# ------------------------------------------------------------------------------------------- long clips in chunks
def _geometry(stft):
    win, hop, fft = stft.window_size, stft.hop_size, stft.fft_size
    pad = stft.pad if stft.center else 0
    return win, hop, fft, pad, -(-win // hop)


def stft_in_chunks(stft, wav, frames_per_chunk, device=None):
    """Generator over (t0, spec[B, 2F, n]) blocks of `stft(wav)` for a clip that need not fit on the device: `wav`
    [B, L] may live on the host (fp32 or int16); every block is computed from the samples its frames touch
    (frames_per_chunk * hop + win - hop of them), copied to `device` on their own.  Concatenated along t the blocks
    equal the whole-clip call to fp32 rounding: each frame is the same transform of the same samples (its pair
    partner in the FFT kernel may differ); the reflect padding (conv_stft.py:54-55) is materialised only for the first
    and last blocks."""
    win, hop, fft, pad, _ = _geometry(stft)
    dev = torch.device(device) if device is not None else stft.weight.device
    if wav.dim() == 1:
        wav = wav[None]
    B, L = wav.shape
    if pad >= L:
        raise RuntimeError(f'stft_in_chunks: reflect padding ({pad}) must be smaller than the clip ({L})')
    T = (L + 2 * pad - win) // hop + 1
    old_center = stft.center
    try:
        stft.center = False                                   # blocks carry their own (real or mirrored) context
        for t0 in range(0, T, frames_per_chunk):
            t1 = min(T, t0 + frames_per_chunk)
            a, b = t0 * hop - pad, (t1 - 1) * hop + win - pad    # sample range in clip coordinates
            core = wav[:, max(a, 0):min(b, L)].to(dev, non_blocking=True)
            parts = [core]
            if a < 0:                                          # F.pad(mode='reflect') of the clip's head
                parts.insert(0, wav[:, 1:1 - a].to(dev).flip(1))
            if b > L:
                parts.append(wav[:, L - 1 - (b - L):L - 1].to(dev).flip(1))
            x = torch.cat(parts, 1) if len(parts) > 1 else core
            with torch.no_grad():
                yield t0, stft(x.contiguous())
    finally:
        stft.center = old_center


def istft_in_chunks(istft, blocks, total_frames, out=None, output_length=None):
    """Overlap-add of a spectrogram delivered as consecutive (t0, spec[B, 2F, n]) blocks (e.g. from
    `stft_in_chunks`, possibly processed in between) into `out` [B, L'] (host or device; a pinned host tensor is
    allocated when None).  Block k owns the samples [t0 * hop, t1 * hop) (the last block: up to the end): they are
    reached only by its own frames and by the J - 1 = ceil(win / hop) - 1 frames before it, which are carried over
    from the previous block and synthesised again -- the overlap-save form of conv_stft.py:101-114.  The same frames
    enter every sum in the same order with the same normaliser, so the result equals `istft(whole spectrogram)` to
    fp32 rounding (frames are transformed in pairs, and a frame's partner depends on where its block starts)."""
    win, hop, fft, pad, J = _geometry(istft)
    if win < hop:
        raise RuntimeError('istft_in_chunks: window shorter than the hop')
    T = total_frames
    full = (T - 1) * hop + win
    start = pad
    end = full - pad if (istft.center and output_length is None) else full
    if output_length is not None:
        end = min(end, start + output_length)
    n_out = max(end - start, 0)
    tail = None                                                # the last <= J - 1 frames seen so far
    for t0, blk in blocks:
        if out is None:
            out = torch.empty((blk.shape[0], n_out),
                              dtype=torch.int16 if istft.pcm16_out else torch.float32).pin_memory()
        t1 = t0 + blk.shape[2]
        nl = tail.shape[2] if tail is not None else 0
        spec = torch.cat([tail, blk], 2).contiguous() if nl else blk.contiguous()
        lo, hi = (t0 * hop if t0 > 0 else 0), (full if t1 >= T else t1 * hop)
        lo_c, hi_c = max(lo, start), min(hi, end)
        if hi_c > lo_c:
            base = (t0 - nl) * hop                             # clip coordinate of the local buffer's first sample
            with torch.no_grad():
                wav = istft._synthesize(spec, lo_c - base, hi_c - lo_c)
            out[:, lo_c - start:hi_c - start].copy_(wav, non_blocking=True)
        if J > 1:
            tail = spec[:, :, -(J - 1):]
    if out is not None and out.device.type == 'cpu':
        torch.cuda.current_stream(istft.weight.device).synchronize()
    return out


def roundtrip_in_chunks(stft, istft, wav, frames_per_chunk, fn=None, device=None):
    """istft(fn(stft(wav))) for a clip of any length, block by block (`fn` maps a spectrogram block [B, 2F, n] to a
    block of the same shape, frame-locally; None = identity).  The ConvSTFT -> ConviSTFT round trip of BASELINE
    config 5 at clip lengths beyond device memory; returns a pinned host tensor [B, L']."""
    if wav.dim() == 1:
        wav = wav[None]
    win, hop, fft, pad, J = _geometry(stft)
    T = (wav.shape[1] + 2 * pad - win) // hop + 1

    def blocks():
        for t0, spec in stft_in_chunks(stft, wav, frames_per_chunk, device):
            yield t0, (fn(spec) if fn is not None else spec)

    return istft_in_chunks(istft, blocks(), T)


def enhance_long(model, wav, chunk_s=30.0, context_s=2.0, sample_rate=16000, device=None):
    """`model(wav)[1]` for a clip longer than one forward should hold, in chunks of `chunk_s` seconds with
    `context_s` seconds of extra signal on both sides of every chunk; the context's output is discarded.  Chunk
    boundaries sit on hop multiples, so a frame-local model gives the whole-clip result exactly; the recurrent
    models (DCCRN / FRCRN LSTM, causal convolutions) see `context_s` seconds of history instead of all of it.
    `wav` [B, L] fp32 (or int16) on the host or the device; returns a pinned host tensor of the same sample type."""
    stft = model.stft
    hop = stft.hop_size
    dev = torch.device(device) if device is not None else next(model.parameters()).device
    if wav.dim() == 1:
        wav = wav[None]
    B, L = wav.shape
    chunk = max(int(chunk_s * sample_rate) // hop, 1) * hop
    ctx = (int(context_s * sample_rate) + hop - 1) // hop * hop
    pipe = HostPipeline(model, dev, dtype=wav.dtype)
    out = None
    for a in range(0, L, chunk):
        b = min(L, a + chunk)
        lo, hi = max(0, a - ctx), min(L, b + ctx)
        y = pipe._forward(wav[:, lo:hi].to(dev, non_blocking=True).contiguous())
        if out is None:
            # the models return L' <= L samples (conv_stft.py:109-111 crop); keep the whole-clip convention
            out = torch.zeros((B, L), dtype=y.dtype).pin_memory()
        n = min(b - a, y.shape[1] - (a - lo))
        out[:, a:a + n].copy_(y[:, a - lo:a - lo + n], non_blocking=True)
    torch.cuda.current_stream(dev).synchronize()
    return out
