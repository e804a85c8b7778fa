"""Tiny end-to-end run for compute-sanitizer (memcheck): B=2 mixed-size images through the fused bf16 path at 150x150 and 299x299,
plus the stand-alone GEMM entry and the static-scratch preprocessing entry."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import kdl_b200
from kdl_b200 import _lib, engine, weights
W = weights.make_weights(1, "calibrated")
rng = np.random.default_rng(0)
imgs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in [(77, 130), (300, 298)]]
for size in (150, 299):
    ref = None
    # default kernels, then every opt-in kernel variant (tensor-core conv1 is the default on the uint8 route; conv1_tc = 0 makes the
    # uint8 and float routes bit-identical)
    for tuning in ({"conv1_tc": 0}, {}, {"fused_midflow": 1}, {"fused_midflow": 2}, {"fused_midflow": 3}, {"dw_tma_store": 1}, {"fused_sepconv": 0}):
        m = engine.Model(W, max_batch=2, input_size=size, precision="bf16", use_graphs=False, tuning=tuning)
        y = m.predict_uint8(imgs)
        if tuning == {"conv1_tc": 0}:
            x = m.preprocess_uint8(imgs)
            y2 = m.forward_device(x).cpu().numpy()
            assert np.array_equal(y, y2)
        elif ref is None:
            ref = y
        else:
            assert np.array_equal(y, ref), tuning
        m.close()
A = torch.randn((300, 728), device="cuda").to(torch.bfloat16); Wt = torch.randn((728, 728), device="cuda").to(torch.bfloat16)
b = torch.zeros(728, device="cuda")
engine.gemm_bf16(A, Wt, b, None, True)
torch.cuda.synchronize()
lib = _lib.load()
pk = engine.PackedImages(imgs, pinned=False); d = pk.host.cuda(); out = torch.empty((2, 299, 299, 3), device="cuda")
err = ctypes.create_string_buffer(256)
assert lib.kdl_preprocess_u8_static(d.data_ptr(), *pk.meta_ptrs(), 2, 299, out.data_ptr(), None, err, 256) == 0
torch.cuda.synchronize()
print("sanitize_small ok")
