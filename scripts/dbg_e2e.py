import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
from test_e2e_gpu import _noise_hook
dec = AudioDecoder(seed=0)
out = {}
z = np.load("tests/golden/e2e_N30_P10.npz")
for tag, hook in (("inj", _noise_hook(int(z["seed"]), z["hift_frames"])), ("dev", None)):
    dec.nsf_noise_hook = hook
    wav, mel = dec.token2wav(torch.from_numpy(z["token"]), "t", prompt_token=torch.from_numpy(z["prompt_token"]),
                             prompt_feat=torch.from_numpy(z["prompt_feat"]), embedding=torch.from_numpy(z["embedding"]))
    out["e2e_wav_" + tag] = wav.cpu().numpy(); out["e2e_mel_" + tag] = mel.cpu().numpy()
# HiFT on the reference mel (injected noise) -> isolates the vocoder
dec.nsf_noise_hook = _noise_hook(int(z["seed"]), z["hift_frames"])
w, _ = dec.hift.inference(speech_feat=torch.from_numpy(z["mel"]).unsqueeze(0))
out["e2e_wav_refmel"] = w.cpu().numpy()
z = np.load("tests/golden/stream_N33_P7.npz")
dec.nsf_noise_hook = _noise_hook(int(z["seed"]), z["hift_frames"])
wav = dec.stream_inference(torch.from_numpy(z["token"]), prompt_token=torch.from_numpy(z["prompt_token"]),
                           prompt_feat=torch.from_numpy(z["prompt_feat"]), embedding=torch.from_numpy(z["embedding"]),
                           block_size=int(z["block_size"]), max_token_len=int(z["max_token_len"]))
out["stream_wav_inj"] = wav.numpy()
np.savez_compressed("gpurun_out/dbg_e2e.npz", **out)
print("ok")
