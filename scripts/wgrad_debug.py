import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from speech_enhancement_b200 import _abi
lib = _abi.lib()
dev = 'cuda:0'
def run(B, Cu, Hu, Wu, Cv, Hv, Wv, kh, kw, sh, sw, ph, pw):
    u = torch.randn(B, Cu, Hu, Wu, device=dev)
    v = torch.randn(B, Cv, Hv, Wv, device=dev)
    ub = torch.empty(lib.se_nc2_elems(B, Cu, Hu, Wu), device=dev, dtype=torch.bfloat16)
    _abi.check(lib.se_nc2_from_nchw(_abi.ptr(u), B, Cu, Hu, Wu, _abi.ptr(ub), _abi.stream_ptr()))
    vb = torch.empty(lib.se_nc2_gather_elems(B, Cv, Hv, Wu, kw), device=dev, dtype=torch.bfloat16)
    _abi.check(lib.se_nc2_gather_w(_abi.ptr(v), B, Cv, Hv, Wv, Wu, kw, sw, pw, _abi.ptr(vb), _abi.stream_ptr()))
    torch.cuda.synchronize(); print('nc2 ok', flush=True)
    G = torch.zeros((kh * kw, Cu, Cv), device=dev)
    _abi.check(lib.se_cconv_wgrad(_abi.ptr(ub), Cu, Hu, Wu, _abi.ptr(vb), Cv, Hv, B, kh, kw, sh, ph, _abi.ptr(G), _abi.stream_ptr()))
    torch.cuda.synchronize(); print('wgrad ok', flush=True)
    # reference: G[tap][cu][cv] = sum u[b,cu,hu,wu] v[b,cv,hu*sh+dh-ph, wu*sw+dw-pw]
    vp = torch.nn.functional.pad(v.double(), (pw, pw + kw + sw * Wu, ph, ph + kh + sh * Hu))
    ref = torch.zeros_like(G, dtype=torch.double)
    for dh in range(kh):
        for dw in range(kw):
            vs = vp[:, :, dh:dh + sh * Hu:sh, dw:dw + sw * Wu:sw]
            ref[dh * kw + dw] = torch.einsum('bchw,bdhw->cd', u.double(), vs)
    err = ((G.double() - ref).norm() / ref.norm()).item()
    print(f'Cu={Cu} Cv={Cv} k={kh}x{kw} s={sh}x{sw}: rel err {err:.2e}', flush=True)
cases = [(2, 128, 8, 64, 128, 16, 64, 5, 1, 2, 1, 2, 0), (2, 128, 8, 64, 128, 8, 64, 1, 1, 1, 1, 0, 0), (2, 128, 8, 64, 128, 8, 72, 1, 2, 1, 1, 0, 0),
         (2, 128, 8, 64, 128, 16, 65, 5, 2, 2, 1, 2, 0), (2, 16, 8, 12, 8, 16, 13, 5, 2, 2, 1, 2, 0),
         (3, 256, 4, 70, 512, 4, 70, 3, 3, 1, 1, 1, 1), (2, 64, 9, 40, 32, 18, 80, 5, 3, 2, 2, 2, 1)]
for c in cases[int(sys.argv[1]):]:
    run(*c)
