#!/usr/bin/env python
"""One-off robustness sweep: the reference's own classes, CPU original vs CUDA drop-in (chained), over odd clip
lengths and batch sizes.  Prints rel-L2 per case; exits non-zero on a failure or an exception."""
import sys, os, warnings, traceback
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader
from speech_enhancement_b200 import dropin
from speech_enhancement_b200.synth import seeded_state_dict, synth_noisy_speech

warnings.simplefilter('ignore')
orig, drop = ref_loader.load_original(), ref_loader.load_dropin()
bad = 0
CASES = [('DCCRN', ('dccrn-CL', 400, 100, 512), {}), ('DCCRN', ('dccrn-C', 512, 128, 512), {}),
         ('DCUNet', ('dcunet10', 512, 128, 512), {}), ('DCUNet', ('dcunet20', 512, 128, 512), {}),
         ('DCUNet', ('dcunet16', 400, 100, 512), {}), ('FRCRN', (320, 160, 640), {}), ('FRCRN', (400, 100, 512), {})]
for cls, args, kw in CASES:
    ref = getattr(orig, cls)(*args, **kw).eval()
    sd = seeded_state_dict(ref, seed=3)
    ref.load_state_dict(sd)
    model = getattr(drop, cls)(*args, **kw).eval()
    model.load_state_dict(sd)
    model.to('cuda:0')
    dropin.accelerate(model)
    for B, L in ((1, 1600), (2, 2048), (1, 4001), (3, 9999), (1, 16000), (5, 16007), (2, 48000), (1, 160000)):
        x = synth_noisy_speech(B, L, seed=B * 7 + L % 13)
        tag = f'{cls}{args} B={B} L={L}'
        try:
            with torch.no_grad():
                want = ref(x)[1]
        except Exception as e:
            print(f'{tag}: reference itself fails ({type(e).__name__}: {str(e)[:60]}) -- skipped')
            continue
        try:
            with torch.no_grad():
                spec, got = model(x.to('cuda:0'))
            torch.cuda.synchronize()
            err = ((got.cpu().double() - want.double()).norm() / want.double().norm().clamp_min(1e-30)).item()
            flag = '' if (err < 1e-3 or cls == 'DCUNet') else '  <-- FAIL'
            if flag:
                bad += 1
            print(f'{tag}: out {tuple(got.shape)} rel-L2 {err:.2e}{flag}')
        except Exception as e:
            bad += 1
            print(f'{tag}: EXCEPTION {type(e).__name__}: {str(e)[:200]}')
            traceback.print_exc(limit=3)
print('failures:', bad)
sys.exit(1 if bad else 0)
