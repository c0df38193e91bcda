import sys, os, numpy as np, torch
sys.path.insert(0, '/root/repo')
from xtts_b200 import _cabi
if len(sys.argv) > 1: _cabi.LIB_PATH = sys.argv[1]
from oracle import mel_oracle as orc
from xtts_b200 import TorchMelSpectrogram
dev = torch.device('cuda:0')
g = np.load('/root/repo/tests/golden/v1_signals.npz')
m = TorchMelSpectrogram(mel_norm_file=None); m.mel_norms = torch.from_numpy(g['norms'])
x = g['x_tone']
y = m(torch.from_numpy(x).to(dev)).cpu().numpy().astype(np.float64)
truth = orc.restated_logmel(x, orc.V1, g['norms'])
ref = g['y_tone'].astype(np.float64)
bound = np.maximum(1e-4, 1.5 * np.abs(ref - truth))
r = np.abs(y - truth) / bound
i = np.unravel_index(r.argmax(), r.shape)
print(sys.argv[1:] , 'worst ratio', r.max(), 'at', i, 'ours err', abs(y[i]-truth[i]), 'ref err', abs(ref[i]-truth[i]), 'value', truth[i], 'n>1', int((r>1).sum()))
for kind in ('puretone','tone_silence','pink'):
    xx = orc.synth(kind, 2, 22050, 5)
    yy = m(torch.from_numpy(xx).to(dev)).cpu().numpy().astype(np.float64)
    tt = orc.restated_logmel(xx, orc.V1, g['norms'])
    rr = orc.callsite_logmel(xx, orc.V1, g['norms']).numpy().astype(np.float64)
    bb = np.maximum(1e-4, 1.5*np.abs(rr-tt))
    q = np.abs(yy-tt)/bb
    print(kind, 'worst ratio', q.max(), 'ours max err', np.abs(yy-tt).max(), 'ref max err', np.abs(rr-tt).max())
