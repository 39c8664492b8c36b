"""SURVEY.md 8f rank 4 -- the wire side of the hot path, on the GPU through the C ABI:
PCM_S 16 samples read by the STFT's stage load and written by the iSTFT's overlap-add epilogue, the int16
HostPipeline, and clips longer than one call processed in chunks (overlap-save)."""
import pytest
import torch

import speech_enhancement_b200 as se
from speech_enhancement_b200 import callers, functional as SF, pipeline
from speech_enhancement_b200.synth import seeded_state_dict, synth_noisy_speech
from conftest import rand, rel_l2

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _pcm(batch, length, seed, scale=0.3):
    x = (rand((batch, length), seed) * scale).clamp(-1.2, 1.2)
    return torch.clamp((x * 32768).round(), -32768, 32767).to(torch.int16)


def _saver_rule(wav):
    """what the reference's savers make of an fp32 waveform (mix_audio.py:144-146, trainer.py:255-262): clamp(-1, 1),
    FFmpeg flt -> s16 = av_clip_int16(lrintf(x * 2^15)) (round half to even, like torch.round)"""
    return torch.clamp((torch.clamp(wav, -1, 1) * 32768).round(), -32768, 32767).to(torch.int16)


@pytest.mark.parametrize('L', [20011, 20111, 1000])        # even / odd frame counts, a clip shorter than one tile
@pytest.mark.parametrize('win,hop,fft', [(400, 100, 512), (512, 128, 512), (320, 160, 512), (320, 160, 640)])
def test_stft_reads_pcm16_and_istft_writes_pcm16_bit_exact(win, hop, fft, L, built_lib):
    st, ist = se.ConvSTFT(win, hop, fft).to(DEV), se.ConviSTFT(win, hop, fft).to(DEV)
    x16 = _pcm(3, L, 5).to(DEV)
    xf = x16.float() / 32768                                  # torchaudio.load's normalisation
    with torch.no_grad():
        n0 = built_lib.se_launch_count()
        spec16 = st(x16)
        fused = built_lib.se_launch_count() - n0 == 1          # one launch: no separate conversion pass
        spec = st(xf)
        assert torch.equal(spec16, spec)                       # the same samples enter the same arithmetic
        assert fused == (built_lib.se_stft_fft_supported(win, hop, fft) == 1)
        big = spec * 4.0                                       # drives part of the waveform beyond +-1: clamp + saturation
        want = _saver_rule(ist(big))
        ist.pcm16_out = True
        n0 = built_lib.se_launch_count()
        got = ist(big)
        assert (built_lib.se_launch_count() - n0 == 1) == fused
    assert got.dtype == torch.int16 and got.shape == want.shape
    assert torch.equal(got, want)
    if L > 5000:
        assert (got == 32767).any() and (got == -32768).any()  # the saturated ends were exercised
    assert torch.equal(SF.pcm16_from_float(xf * 1.5), _saver_rule(xf * 1.5))
    assert torch.equal(SF.pcm16_to_float(x16), xf)


def test_pcm16_entry_points_reject_bad_arguments(built_lib):
    st = se.ConvSTFT(400, 100, 512).to(DEV)
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        st(torch.zeros(2, 4000, dtype=torch.int16))
    ist = se.ConviSTFT(400, 100, 512).to(DEV)
    ist.pcm16_out = True
    spec = torch.zeros(1, 514, 20, device=DEV, requires_grad=True)
    with pytest.raises(RuntimeError, match='inference option'):
        ist(spec)
    assert built_lib.se_pcm16_from_f32(None, 4, None, None) == -1
    assert built_lib.se_pcm16_from_f32(None, 0, None, None) == 0      # nothing to convert: no pointers needed


def test_host_pipeline_pcm16_matches_fp32_pipeline_and_halves_the_copies():
    model = callers.DCCRN('dccrn-CL', 400, 100, 512).eval()
    model.load_state_dict(seeded_state_dict(model, seed=3))
    model.to(DEV)
    B, L = 3, 16000
    x16 = torch.clamp((synth_noisy_speech(B, L, seed=8) * 32768).round(), -32768, 32767).to(torch.int16)
    xf = (x16.float() / 32768)
    p32, p16 = pipeline.HostPipeline(model, DEV), pipeline.HostPipeline(model, DEV, dtype=torch.int16)
    n = p32.output_length(xf)
    assert p16.output_length(x16) == n
    out32 = [torch.empty((B, n)).pin_memory() for _ in range(2)]
    out16 = [torch.empty((B, n), dtype=torch.int16).pin_memory() for _ in range(2)]
    for i in range(3):                                         # slots are reused from the third submit on
        p32.submit(xf.pin_memory(), out32[i & 1])
        p16.submit(x16.pin_memory(), out16[i & 1])
    p32.synchronize()
    p16.synchronize()
    assert torch.equal(out16[0], _saver_rule(out32[0])) and torch.equal(out16[1], _saver_rule(out32[1]))
    assert not model.istft.pcm16_out                           # the pipeline restores the module switch
    assert x16.element_size() * 2 == xf.element_size()
    with pytest.raises(RuntimeError, match='must be torch.int16'):
        p16.submit(xf.pin_memory(), out16[0])


def test_host_pipeline_pcm16_through_the_accelerated_reference_class(built_lib):
    """the int16 pipeline must take the CHAINED forward of a dropin.accelerate()d reference model (the reference's own
    forward would clamp_ the int16 samples to [-1, 1])"""
    from oracle import ref_loader
    from speech_enhancement_b200 import dropin
    if not ref_loader.available():
        pytest.skip(ref_loader.why_unavailable())
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        model = ref_loader.load_dropin().DCCRN('dccrn-CL', 400, 100, 512).eval()
        model.load_state_dict(seeded_state_dict(model, seed=3))
        model.to(DEV)
    dropin.accelerate(model)
    B, L = 2, 16000
    x16 = torch.clamp((synth_noisy_speech(B, L, seed=8) * 32768).round(), -32768, 32767).to(torch.int16)
    p16 = pipeline.HostPipeline(model, DEV, dtype=torch.int16)
    n = p16.output_length(x16)
    out = torch.empty((B, n), dtype=torch.int16).pin_memory()
    calls = []
    view = model._se_chained
    inner = view._forward_fused
    view.__dict__['_forward_fused'] = lambda inp: (calls.append(inp.dtype), inner(inp))[1]
    p16.submit(x16.pin_memory(), out)
    p16.synchronize()
    assert calls == [torch.int16]
    with torch.no_grad():
        want = _saver_rule(model((x16.float() / 32768).to(DEV))[1]).cpu()
    assert torch.equal(out, want)
    # outside the chained path an int16 waveform must not reach the reference's clamp_
    model.istft.pcm16_out = True
    model.train()
    with pytest.raises(RuntimeError, match='pcm16_out needs the chained forward'), warnings.catch_warnings():
        warnings.simplefilter('ignore')
        model(x16.to(DEV))
    model.istft.pcm16_out = False


@pytest.mark.parametrize('win,hop,fft,frames', [(400, 100, 512, 1000), (512, 128, 512, 777), (320, 160, 640, 500)])
def test_60_s_clip_in_chunks_equals_the_whole_clip(win, hop, fft, frames):
    st, ist = se.ConvSTFT(win, hop, fft).to(DEV), se.ConviSTFT(win, hop, fft).to(DEV)
    L = 60 * 16000
    wav = synth_noisy_speech(2, L, seed=13)                     # stays on the host; blocks travel one by one
    with torch.no_grad():
        spec = st(wav.to(DEV))
        whole = ist(spec)
    T = spec.shape[2]
    blocks = list(pipeline.stft_in_chunks(st, wav, frames, DEV))
    assert [b[0] for b in blocks] == list(range(0, T, frames))
    assert rel_l2(torch.cat([b[1] for b in blocks], 2).cpu(), spec.cpu()) <= 1e-6
    assert st.center                                           # restored
    got = pipeline.istft_in_chunks(ist, iter(blocks), T)
    assert got.shape == whole.shape and got.is_pinned()
    assert rel_l2(got, whole.cpu()) <= 1e-6
    assert (got - whole.cpu()).abs().max().item() <= 1e-5
    rt = pipeline.roundtrip_in_chunks(st, ist, wav, frames, fn=lambda s: s * 0.5)
    assert rel_l2(rt, 0.5 * whole.cpu()) <= 1e-6
    # output_length and PCM16 out through the chunked path
    ist.pcm16_out = True
    with torch.no_grad():
        whole16 = ist(spec, output_length=L - 123)
    got16 = pipeline.istft_in_chunks(ist, iter(blocks), T, output_length=L - 123)
    assert got16.dtype == torch.int16 and got16.shape == whole16.shape
    assert (got16.int() - whole16.cpu().int()).abs().max().item() <= 1   # fp32 rounding may move a sample by one LSB


def test_pcm16_empty_batch_and_recurrent_model_in_chunks():
    st, ist = se.ConvSTFT(400, 100, 512).to(DEV), se.ConviSTFT(400, 100, 512).to(DEV)
    ist.pcm16_out = True
    with torch.no_grad():
        spec = st(torch.zeros((0, 4000), dtype=torch.int16, device=DEV))
        assert spec.shape == (0, 514, 42)
        out = ist(spec)
        assert out.shape == (0, 3988) and out.dtype == torch.int16
    # a recurrent model through enhance_long: context makes it an approximation, but it must run, keep the length
    # convention and agree with the whole-clip result away from the start of every chunk
    model = callers.DCCRN('dccrn-CL', 400, 100, 512).eval()
    model.load_state_dict(seeded_state_dict(model, seed=3))
    model.to(DEV)
    wav = synth_noisy_speech(1, 6 * 16000, seed=21)
    with torch.no_grad():
        whole = model(wav.to(DEV))[1].cpu()
    got = pipeline.enhance_long(model, wav, chunk_s=2.0, context_s=1.0)
    assert got.shape == (1, 6 * 16000) and torch.isfinite(got).all()
    n = whole.shape[1]
    assert rel_l2(got[:, :16000], whole[:, :16000]) < 1e-4           # the first chunk sees the same history
    assert not model.istft.pcm16_out


class _FrameLocal(torch.nn.Module):
    """a caller whose network is frame-local (a fixed spectral gain): chunked == whole clip"""

    def __init__(self):
        super().__init__()
        self.stft, self.istft = se.ConvSTFT(400, 100, 512), se.ConviSTFT(400, 100, 512)
        self.gain = torch.nn.Parameter(torch.linspace(0.2, 1.0, 514)[None, :, None])

    def forward(self, x):
        spec = self.stft(x) * self.gain
        wav = self.istft(spec)
        return spec, wav if wav.dtype == torch.int16 else torch.clamp_(wav, -1, 1)


def test_enhance_long_matches_whole_clip_for_a_frame_local_model():
    model = _FrameLocal().to(DEV).eval()
    L = 20 * 16000
    wav = synth_noisy_speech(2, L, seed=17)
    with torch.no_grad():
        whole = model(wav.to(DEV))[1].cpu()
    got = pipeline.enhance_long(model, wav, chunk_s=3.0, context_s=0.5)
    assert got.shape == (2, L)
    n = whole.shape[1]
    assert rel_l2(got[:, :n], whole) <= 1e-6
    got16 = pipeline.enhance_long(model, torch.clamp((wav * 32768).round(), -32768, 32767).to(torch.int16),
                                  chunk_s=3.0, context_s=0.5)
    assert got16.dtype == torch.int16
    assert (got16[:, :n].int() - _saver_rule(whole).int()).abs().max().item() <= 1
