import numpy as np, torch, sys, os
sys.path.insert(0, os.getcwd())
from moss_speech_decoder_cosy_b200 import runtime
from oracle import rng_replay
z=np.load('tests/golden/config1_N125.npz')
eng=runtime.build_engine(seed=0)
mel=torch.from_numpy(z['mel']).cuda()
noise=rng_replay.replay(int(z['noise_seed']),[480*500])[0].cuda()
wav,src=eng.hift([500],mel.reshape(-1),nsf_noise=noise)
torch.cuda.synchronize()
f0=eng.debug_read("hift.f0",(500+8)*4).reshape(-1,4)[4:504,0]
f0r=torch.from_numpy(z['f0'])
v=f0r>10
d=(f0-f0r)
print('voiced',int(v.sum()),'mean signed abs err',float(d[v].mean()),'mean signed rel',float((d/f0r)[v].mean()),'rms rel',float((d/f0r)[v].pow(2).mean().sqrt()),'max rel voiced',float((d.abs()/f0r)[v].max()))
print('cum drift of fundamental (rad) at end', float((d*v).cumsum(0)[-1]/24000*480*2*np.pi))
