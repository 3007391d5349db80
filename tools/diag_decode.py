"""Diagnostic (GPU box): where does the EXACT-mode decode of the unperturbed theta differ from the oracle?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden'))
import numpy as np, torch
import nes_b200, make_golden
from oracle import fc_caption, noise

for name in sys.argv[1:] or ['g3']:
    c = make_golden.CASES[name]
    dims, G, play, D, D_pad, L, theta, fc, offs = make_golden.case_inputs(c)
    mo = dict(vocab_size=c['V'], input_encoding_size=c['E'], rnn_size=c['R'], fc_feat_size=c['Fd'], layer_n=c.get('ln', False), layer_n_affine=c.get('affine', False))
    ctx = nes_b200.NesContext(mo, seed=make_golden.TABLE_SEED, table_len=L, wave_members=4, rows=c['B'])
    ctx.set_theta(theta)
    s_ref, lp_ref, tr = fc_caption.decode_greedy(dims, theta, fc, trace=True)
    for prec in (1, 0):
        seq, lp, h_tr, l_tr = ctx.rollout(np.zeros(1, dtype=np.int64), np.ones(1, dtype=np.float32), fc, 0.0, precision=prec, trace=True)
        seq = seq[0].numpy(); l_tr = l_tr.cpu().numpy()[:, 0]; h_tr = h_tr.cpu().numpy()[:, 0]
        bad = np.argwhere(seq != s_ref)
        print(name, 'prec', prec, 'mismatches', len(bad), 'of', seq.size, 'steps run by oracle', len(tr['logits']))
        if len(bad):
            r, t = bad[np.argmin(bad[:, 1])]
            lo = tr['logits'][t][r]; top = np.argsort(lo)[-3:][::-1]
            print('  first mismatch row %d step %d: gpu tok %d ref tok %d; ref top3 %s logits %s; gpu logits there %s' % (r, t, seq[r, t], s_ref[r, t], top, lo[top], l_tr[t, r, top]))
            print('  max |logit err| at that step/row', np.abs(l_tr[t, r] - lo).max(), 'max|logit|', np.abs(lo).max())
        for t in (0, 1, 8, 16):
            if t < len(tr['h']):
                print('  h err t=%d: %.3e (max|h| %.3e)' % (t, np.abs(h_tr[t] - tr['h'][t]).max(), np.abs(tr['h'][t]).max()))
        for t in (0, 7, 15):
            if t < len(tr['logits']):
                print('  logit err t=%d: %.3e (max|l| %.3e)' % (t, np.abs(l_tr[t] - tr['logits'][t]).max(), np.abs(tr['logits'][t]).max()))
    ctx.close()
