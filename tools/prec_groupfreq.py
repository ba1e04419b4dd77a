import numpy as np, torch, sys
sys.path.insert(0, '.')
from pydsproutines_b200 import signals
import pydsproutines_b200.xcorrRoutines as xr
from oracle import xcorr_oracle as orc
c = signals.config_c3(log2_shifts=14)
F = c["F"]; NFQ = 256
freqs = (np.arange(NFQ) - 100) / F + c["k0"] / F - 28.0 / F
y = np.zeros(int(c["starts"][-1]) + 128, np.complex64)
for g in range(64): y[int(c["starts"][g]):int(c["starts"][g]) + 128] = c["ygroups"][g]
a = xr.GroupXcorr(y, c["starts"], np.full(64, 128), freqs, 1.0)
K = c["K"]
xc, fp = a.xcorr(torch.from_numpy(c["rx"]).cuda(), np.arange(K))
xc = xc.cpu().numpy()
o = orc.GroupXcorr(y, c["starts"], np.full(64, 128), freqs, 1.0)
rng = np.random.default_rng(1)
sample = np.unique(np.concatenate([rng.integers(0, K, 60), [c["p0"]]]))
rxc, _ = o.xcorr(c["rx"], sample)
rel = (xc[sample] - rxc) / rxc
print("peak rel err %.3e ; off-peak rel err mean %.3e max|.| %.3e" % (rel[sample == c["p0"]][0], rel.mean(), np.abs(rel).max()))
