"""How many vocabulary columns fall within the HYBRID margin of the row max?  (fp16-operand logit trace, bench-like setup)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import nes_b200
V, E, R, FD = 9487, 512, 512, 2048
P, rows = 4, 128
ctx = nes_b200.NesContext(dict(vocab_size=V, input_encoding_size=E, rnn_size=R, fc_feat_size=FD), seed=1234, table_len=1 << 26, wave_members=P, rows=rows)
model = nes_b200.FCModel(ctx, parent=0, rng_state=0)
fc = torch.randn(rows, FD, generator=torch.Generator().manual_seed(0))
offs = ctx.sample_offsets(0, P)
sigma = 0.005
seq, lp, h_tr, l_tr = ctx.rollout(offs, torch.ones(P), fc, sigma, precision=nes_b200.PREC_FAST, trace=True)
h = h_tr[1:].double()                      # [T, P, B, R] state feeding the logits of step t
lt = l_tr.double()                         # [T, P, B, V1]
theta = ctx.get_theta()
lay = {t.name: t for t in ctx.tensors}
W = theta[lay['logit.weight'].start:lay['logit.weight'].start + lay['logit.weight'].count].view(V + 1, R).double()
th_max = W.norm(dim=1).max().item()
tab = torch.from_numpy(ctx.table_slice(0, 1 << 24).astype(np.float32)).view(-1, R)
e_max = tab.norm(dim=1).max().item()
xn = h.norm(dim=-1)                        # [T, P, B]
margin = 2 * 1.1 * 2 ** -11 * xn * (2 * th_max + sigma * e_max)
mx = lt.max(-1).values
cnt = (lt >= (mx - margin).unsqueeze(-1)).sum(-1)          # includes the max itself
spread = lt.std(-1)
print('theta_rowmax %.4f eps_rowmax %.3f |h| mean %.3f margin/spread mean %.4f' % (th_max, e_max, xn.mean(), (margin / spread).mean()))
print('candidates per row: mean %.3f  max %d  rows with > 4: %.5f' % (cnt.double().mean(), cnt.max(), (cnt > 4).double().mean()))
for t in range(16):
    print(t, 'cand mean %.3f max %d  |h| %.3f  spread %.4f  margin %.5f  top gap %.5f' % (cnt[t].double().mean(), cnt[t].max(), xn[t].mean(), spread[t].mean(), margin[t].mean(),
          (lt[t].topk(2, -1).values[..., 0] - lt[t].topk(2, -1).values[..., 1]).mean()))
ctx.profile(True)
seq2, lp2 = ctx.rollout(offs, torch.ones(P), fc, sigma, precision=nes_b200.PREC_HYBRID)
print('kernel stats', ctx.hybrid_stats())
seq1, lp1 = ctx.rollout(offs, torch.ones(P), fc, sigma, precision=nes_b200.PREC_EXACT)
print('tokens equal to EXACT:', bool((seq1 == seq2).all()), 'max |dlogp|', float((lp1 - lp2).abs().max()))
