"""NIC.py dense voxel encoder (SURVEY 8a A1''): forward / backward / apply_gradients timings on one B200 with the
roofline that bounds each (HBM: W read once forward, dW written once backward, 28 B/param in the optimizer)."""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch

pkg = importlib.import_module("masters-thesis_b200")
dev = torch.device("cuda", 0)
K, E, B = 327684, 512, 64
peak = 6547.8
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
enc = pkg.DenseEncoder(K, E, 0.0, 0.2, 0.01, batch_max=B, device=dev)
enc.compile(pkg.Adam(1e-4, beta_1=0.9, beta_2=0.98, epsilon=10.0e-9, clipnorm=0.1))
xs = [torch.randn(B, K, device=dev) for _ in range(2)]
df = torch.randn(B, E, device=dev) / B
def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(n):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
i = [0]
def fwd():
    i[0] += 1
    enc(xs[i[0] % 2], training=True)
def step():
    fwd(); enc.backward(df); enc.apply_gradients()
t_f = timed(fwd)
t_fb = timed(lambda: (fwd(), enc.backward(df)))
t_all = timed(step)
W = K * E * 4
out = {"workload": f"NIC.py dense encoder K={K} E={E} B={B} fp32", "forward_ms": t_f, "backward_ms": t_fb - t_f,
       "apply_gradients_ms": t_all - t_fb, "step_ms": t_all, "samples_per_s": B / (t_all * 1e-3),
       "forward_roofline": {"bound": "hbm", "alg_bytes": W + 2 * B * K * 4 * 1.0, "achieved_GBps": (W + 2 * B * K * 4) / t_f / 1e6,
                            "peak": peak, "frac": (W + 2 * B * K * 4) / t_f / 1e6 / peak,
                            "note": "W read once + x read and xT written; fp32 FFMA ceiling 21.5 GF / 74.4 TF/s = 0.29 ms"},
       "backward_roofline": {"bound": "hbm", "alg_bytes": W + B * K * 4, "achieved_GBps": (W + B * K * 4) / (t_fb - t_f) / 1e6,
                             "peak": peak, "frac": (W + B * K * 4) / (t_fb - t_f) / 1e6 / peak},
       "optimizer_roofline": {"bound": "hbm", "alg_bytes": 7 * W + 2 * W, "achieved_GBps": 9 * W / (t_all - t_fb) / 1e6,
                              "peak": peak, "frac": 9 * W / (t_all - t_fb) / 1e6 / peak,
                              "note": "norm pass reads w,g (8 B/param) + Adam reads w,m,v,g writes w,m,v (28 B/param)"}}
print(json.dumps(out))
