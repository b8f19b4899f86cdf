import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from moss_speech_decoder_cosy_b200 import runtime
eng = runtime.build_engine(seed=0)
out = {}
for name in ["hift_T40", "hift_T28_cache"]:
    z = np.load(f"tests/golden/{name}.npz")
    mel = torch.from_numpy(z["mel"]).cuda(); T = mel.shape[1]
    noise = torch.from_numpy(z["noise"]).cuda(); cache = torch.from_numpy(z["cache"]).cuda()
    kw = dict(cache_source=cache.reshape(1, -1), cache_len=[cache.numel()]) if cache.numel() else {}
    wav, src = eng.hift([T], mel.reshape(-1), nsf_noise=noise, **kw)
    torch.cuda.synchronize()
    out[name + ".wav"] = wav.cpu().numpy(); out[name + ".src"] = src.cpu().numpy()
    out[name + ".f0"] = eng.debug_read("hift.f0", (T + 8) * 4).numpy()
    out[name + ".p480"] = eng.debug_read("hift.p480", (T + 8) * 16).numpy()
    out[name + ".post"] = eng.debug_read("hift.post", (120 * (T + 8)) * 32).numpy()
os.makedirs("gpurun_out", exist_ok=True)
np.savez_compressed("gpurun_out/dbg_hift.npz", **out)
print("ok")
