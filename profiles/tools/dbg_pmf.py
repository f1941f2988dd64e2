import importlib, torch, sys
sys.path.insert(0, '/root/repo')
em = importlib.import_module('l-msfc-v2-training_b200.entropy_models'); mm = importlib.import_module('l-msfc-v2-training_b200.model')
DEV='cuda'
torch.manual_seed(3)
eb_h = em.EntropyBottleneck(8).to(DEV)
eb_d = em.EntropyBottleneck(8).to(DEV); eb_d.load_state_dict(eb_h.state_dict()); eb_d.device_pmf=True
eb_h.update(force=True); eb_d.update(force=True)
d=(eb_h._quantized_cdf-eb_d._quantized_cdf).abs(); print('EB init', int(d.max()), eb_h._quantized_cdf[0].tolist(), eb_d._quantized_cdf[0].tolist())
gc_h = em.GaussianConditional(None).to(DEV); gc_d = em.GaussianConditional(None).to(DEV); gc_d.device_pmf=True
t = mm.get_scale_table(); gc_h.update_scale_table(t, force=True); gc_d.update_scale_table(t, force=True)
d=(gc_h._quantized_cdf-gc_d._quantized_cdf).abs(); print('GC', int(d.max()), (d>0).sum().item(), d.numel())
r=int(d.max(1).values.argmax()); print('row', r, gc_h._quantized_cdf[r,:12].tolist(), gc_d._quantized_cdf[r,:12].tolist(), gc_h._cdf_length[r].item())
L = importlib.import_module('l-msfc-v2-training_b200._lib')
import scipy.stats
table = gc_h.scale_table.detach().cpu()
mult = -scipy.stats.norm.ppf(1e-9 / 2)
center = torch.ceil(table * mult).int(); plen = 2 * center + 1; ml = int(plen.max())
samples = torch.abs(torch.arange(ml).int() - center[:, None]).float(); sc = table.unsqueeze(1).float(); const = float(-(2 ** -0.5))
upper = 0.5 * torch.erfc(const * ((0.5 - samples) / sc)); lower = 0.5 * torch.erfc(const * ((-0.5 - samples) / sc)); pmf_h = upper - lower
pmf_d = torch.empty((64, ml), device=DEV); tail = torch.empty(64, device=DEV)
L.gc_pmf(L.ptr(gc_h.scale_table), L.ptr(center.to(DEV)), 64, ml, L.ptr(pmf_d), ml, L.ptr(tail), L.stream()); torch.cuda.synchronize()
pd = pmf_d.cpu(); r = 41; n = int(plen[r])
a, b = pmf_h[r, :n], pd[r, :n]
ca, cb = torch.round(a * 65536), torch.round(b * 65536)
print('row41 len', n, 'count diffs', int((ca != cb).sum()), 'max rel', float(((a - b).abs() / a.clamp_min(1e-30)).max()))
idx = (ca != cb).nonzero().flatten()[:8].tolist(); print([(i, float(a[i]) * 65536, float(b[i]) * 65536) for i in idx])
print('tail', float(2 * lower[r, 0]), float(tail[r]))
ca_ = gc_h._pmf_to_cdf(pmf_h, 2 * lower[:, :1], plen, ml); cb_ = gc_h._pmf_to_cdf(pmf_d, tail, plen, ml)
print('direct: host-pmf vs device-pmf', int((ca_ - cb_).abs().max()), ' host-pmf vs gc_h', int((ca_ - gc_h._quantized_cdf).abs().max()),
      ' device-pmf vs gc_d', int((cb_ - gc_d._quantized_cdf).abs().max()))
cnt_h, cnt_d = torch.round(pmf_h * 65536), torch.round(pd * 65536)
bad = [(r, int((cnt_h[r, :int(plen[r])] != cnt_d[r, :int(plen[r])]).sum())) for r in range(64)]
print('rows with differing counts', [b for b in bad if b[1]])
print('tails equal', torch.equal((2 * lower[:, 0]), tail.cpu()))
dd = (ca_ - cb_).abs(); rr = int(dd.max(1).values.argmax()); first = int((dd[rr] > 0).nonzero()[0]); n = int(plen[rr])
print('row', rr, 'len', n, 'first diff at', first, ca_[rr, first-2:first+4].tolist(), cb_[rr, first-2:first+4].tolist())
pa, pb = pmf_h[rr, :n], pd[rr, :n]
neq = (pa != pb).nonzero().flatten(); print('raw float differences in row:', neq.numel(), [(int(i), float(pa[i]), float(pb[i])) for i in neq[:6]])
print('tail raw', float(2 * lower[rr, 0]).hex(), float(tail[rr]).hex())
from oracle import cdf as OC
import numpy as np
def oracle_row(p, t, n):
    row = np.concatenate([p[:n].numpy(), np.array([t], np.float32)]).astype(np.float32)
    return np.array(OC.pmf_to_quantized_cdf(row, 16))
oa = oracle_row(pmf_h[rr], float(2 * lower[rr, 0]), n); ob = oracle_row(pd[rr], float(tail[rr]), n)
print('oracle on host pmf == kernel:', np.array_equal(oa, ca_[rr, :n + 2].cpu().numpy()), ' oracle on device pmf == kernel:', np.array_equal(ob, cb_[rr, :n + 2].cpu().numpy()),
      ' oracle host vs device pmf equal:', np.array_equal(oa, ob))
ka = np.round(pmf_h[rr, :n].numpy().astype(np.float32) * np.float32(65536)); kb = np.round(pd[rr, :n].numpy().astype(np.float32) * np.float32(65536))
print('count diffs (numpy)', int((ka != kb).sum()), 'sum', ka.sum(), kb.sum(), 'entries 50..56', ka[50:57], kb[50:57])
