"""Output-gradient sensitivities for the safe mutations SM-G-SUM / SM-G-ABS (src/algorithm/safe_mutations.py:14-139,
Lehman et al.'s safe mutations as the reference adapted them) and the file protocol the workers share them through.

SURVEY 8f-3 assigns the Jacobian to PyTorch autograd: it is computed once per (iteration, parent), not per member, on the
device that holds theta; what the population hot path consumes is the resulting vector (`NesContext.set_sensitivity`,
noise /= sensitivity inside the CUDA kernels).  The differentiable forward below restates
`FCModel.forward_for_sensitivity` (src/captioning/nets.py:22-69) over views of ONE flat theta leaf, so the gradient
arrives already in parameter_vector order (the reference gathers it with `extract_grad`, algorithm/nets.py:131-143).
"""
import logging
import os
import re
import time

import torch
import torch.nn.functional as F


def find_file_with_pattern(pattern, directory):
    """src/algorithm/tools/utils.py:149-153"""
    for file in os.listdir(directory):
        if re.search(pattern, file):
            return file


def _views(theta, tensors):
    return {t.name: theta[t.start:t.start + t.count].view(*t.shape) for t in tensors}


def _core(p, dims, xt, h, c):
    """LSTMCore.forward (src/captioning/nets.py:98-134): plain, LayerNorm (layer_n = 1) or batch norm over the rows (= 2, vbn)"""
    R, ln = dims['rnn_size'], int(dims.get('layer_n') or 0)
    a_i = F.linear(xt, p['core.i2h.weight'], p['core.i2h.bias'])
    a_h = F.linear(h, p['core.h2h.weight'], p['core.h2h.bias'])
    if ln == 1:
        a_i = F.layer_norm(a_i, (5 * R,), p.get('core.i2h_ln.weight'), p.get('core.i2h_ln.bias'), 1e-5)
        a_h = F.layer_norm(a_h, (5 * R,), p.get('core.h2h_ln.weight'), p.get('core.h2h_ln.bias'), 1e-5)
    elif ln == 2:
        a_i = F.batch_norm(a_i, None, None, p.get('core.i2h_bn.weight'), p.get('core.i2h_bn.bias'), True, 0.1, 1e-5)
        a_h = F.batch_norm(a_h, None, None, p.get('core.h2h_bn.weight'), p.get('core.h2h_bn.bias'), True, 0.1, 1e-5)
    s = a_i + a_h
    sig = torch.sigmoid(s.narrow(1, 0, 3 * R))
    in_gate, forget_gate, out_gate = sig.narrow(1, 0, R), sig.narrow(1, R, R), sig.narrow(1, 2 * R, R)
    in_transform = torch.max(s.narrow(1, 3 * R, R), s.narrow(1, 4 * R, R))
    next_c = forget_gate * c + in_gate * in_transform
    if ln == 1:
        act = torch.tanh(F.layer_norm(next_c, (R,), p.get('core.c_ln.weight'), p.get('core.c_ln.bias'), 1e-5))
    elif ln == 2:
        act = torch.tanh(F.batch_norm(next_c, None, None, p.get('core.c_bn.weight'), p.get('core.c_bn.bias'), True, 0.1, 1e-5))
    else:
        act = torch.tanh(next_c)
    return out_gate * act, next_c


def forward_for_sensitivity(theta, tensors, dims, fc_feats, orig_bs=0, i=-1, split=100, length=5, seq_per_img=5):
    """FCModel.forward_for_sensitivity (src/captioning/nets.py:22-69): the log-probs after `length` greedy steps, the vocabulary
    folded into groups of `split` entries by their L2 norm -> [rows, ceil-ish(V1 / split)].  theta: flat float32 tensor
    (requires_grad for the Jacobian); fc_feats: [batch * seq_per_img, Fd], every seq_per_img-th row is used (:41)."""
    p = _views(theta, tensors)
    fc = torch.as_tensor(fc_feats).to(device=theta.device, dtype=theta.dtype)
    fc = fc.index_select(0, torch.arange(0, fc.size(0), seq_per_img, device=fc.device))
    if fc.size(0) > orig_bs > 0:
        fc = fc[:orig_bs]
    if i >= 0:
        fc = fc[i].unsqueeze(0)
    B, R = fc.size(0), dims['rnn_size']
    h = fc.new_zeros(B, R)
    c = fc.new_zeros(B, R)
    vbn_e = bool(dims.get('vbn_e'))                   # Sequential(layer, BatchNorm1d) on both embeddings (captioning/nets.py:156-165)
    n0 = '.0' if vbn_e else ''

    def bn(x, layer):
        return F.batch_norm(x, None, None, p.get(layer + '.1.weight'), p.get(layer + '.1.bias'), True, 0.1, 1e-5) if vbn_e else x
    h, c = _core(p, dims, bn(F.linear(fc, p['img_embed%s.weight' % n0], p['img_embed%s.bias' % n0]), 'img_embed'), h, c)
    it = torch.zeros(B, dtype=torch.long, device=fc.device)
    logprobs = None
    for _ in range(length):
        h, c = _core(p, dims, bn(F.embedding(it, p['embed%s.weight' % n0]), 'embed'), h, c)
        logprobs = F.log_softmax(F.linear(h, p['logit.weight'], p['logit.bias']), dim=1)
        it = torch.max(logprobs.detach(), 1)[1].view(-1).long()
    cat_size = split - (logprobs.size(1) % split)
    extended = torch.cat((logprobs, logprobs.new_zeros(B, cat_size)), 1)
    return ((torch.stack(extended.split(split, dim=1)) ** 2).sum(2) ** (1 / 2)).permute(1, 0)


def mnist_forward(theta, tensors, images):
    """MnistNet.forward (src/classification/nets.py:25-37) over views of one flat theta; images [B, 1, 28, 28] -> logits [B, 10]"""
    p = _views(theta, tensors)
    x = torch.as_tensor(images).to(device=theta.device, dtype=theta.dtype)
    x = F.max_pool2d(F.relu(F.conv2d(x, p['conv1.weight'], p['conv1.bias'])), 2, 2)
    x = F.max_pool2d(F.relu(F.conv2d(x, p['conv2.weight'], p['conv2.bias'])), 2, 2)
    return F.linear(x.reshape(-1, 4 * 4 * 20), p['fc1.weight'], p['fc1.bias'])


def sum_sensitivity_of(theta, fwd):
    """Sensitivity._calc_sum_sensitivity (safe_mutations.py:93-115): row k of the Jacobian is d(sum over the batch of output k)/dtheta,
    sensitivity = sqrt(sum_k J_k^2) / batch.  fwd(theta, i) -> [rows, K] (i = -1: the whole batch); one backward pass per output."""
    theta = theta.detach().clone().requires_grad_(True)
    with torch.enable_grad(), torch.backends.cudnn.flags(allow_tf32=False):       # float32 convolutions like the reference's CPU run
        out = fwd(theta, -1)
        K, B = out.size(1), out.size(0)
        jac = torch.zeros(K, theta.numel(), dtype=theta.dtype, device=theta.device)
        grad_output = torch.zeros_like(out)
        for k in range(K):
            grad_output.zero_()
            grad_output[:, k] = 1.0
            jac[k] = torch.autograd.grad(out, theta, grad_outputs=grad_output, retain_graph=True)[0]
    sens = torch.sqrt((jac ** 2).sum(0))
    sens /= B
    return sens.detach()


def abs_sensitivity_of(theta, fwd):
    """Sensitivity._calc_abs_sensitivity (safe_mutations.py:119-147) as it is evidently meant: per-sample Jacobians, |.| averaged over
    the batch, then the same root of squares.  UNPINNED: the reference's own method raises AttributeError (`self.nb_params` at :125
    does not exist on Sensitivity), so there is nothing to compare against; one forward + backward per (output, sample)."""
    theta = theta.detach().clone().requires_grad_(True)
    with torch.enable_grad(), torch.backends.cudnn.flags(allow_tf32=False):
        full = fwd(theta, -1)
        K, B = full.size(1), full.size(0)
        acc = torch.zeros(K, theta.numel(), dtype=theta.dtype, device=theta.device)
        for i in range(B):
            out_i = fwd(theta, i)
            grad_output = torch.zeros_like(out_i)
            for k in range(K):
                grad_output.zero_()
                grad_output[0, k] = 1.0
                acc[k] += torch.autograd.grad(out_i, theta, grad_outputs=grad_output, retain_graph=True)[0].abs()
    return torch.sqrt(((acc / B) ** 2).sum(0)).detach()


def sum_sensitivity(theta, tensors, dims, fc_feats, orig_bs=0):
    """SM-G-SUM for fc_caption: sum_sensitivity_of over forward_for_sensitivity"""
    return sum_sensitivity_of(theta, lambda th, i: forward_for_sensitivity(th, tensors, dims, fc_feats, orig_bs, i=i))


def abs_sensitivity(theta, tensors, dims, fc_feats, orig_bs=0):
    """SM-G-ABS for fc_caption (unpinned, see abs_sensitivity_of)"""
    return abs_sensitivity_of(theta, lambda th, i: forward_for_sensitivity(th, tensors, dims, fc_feats, orig_bs, i=i))


class Sensitivity:
    """Per-parent sensitivity vector with the reference's sharing protocol (src/algorithm/safe_mutations.py:14-91): a vector is valid
    for one (task_id, parent_id); the first worker that needs it computes it and leaves `sens_t{task}_p{parent}_{pid}.pt` in
    `directory`, everybody else loads that file.  `net` supplies parameter_vector(), sensitivity_forward() and ctx.dev."""

    FILE = 'sens_t{t}_p{p}_{i}.pt'

    def __init__(self, net, underflow, method):
        self.net, self._underflow, self._type = net, underflow, method
        self._vector = None
        self._valid_for = None              # (task_id, parent_id) the vector belongs to
        self._orig_batch_size = 0           # batch size of the first call; later calls keep using it (:57-58)

    def get_sensitivity(self):
        return self._vector

    def set_sensitivity(self, path):
        """:27-32 -- a precomputed vector (SM-VECTOR), from a torch-saved file or in memory: floor at the underflow value, then
        scale so that the smallest entry is 1"""
        vec = torch.load(path, map_location='cpu') if isinstance(path, str) else torch.as_tensor(path).clone()
        vec = vec.to(torch.float32).clamp_(min=self._underflow)
        self._vector = vec / vec.min()

    def calc_sensitivity(self, task_id, parent_id, experiences, batch_size, directory):
        """:34-91.  Returns True when the vector changed (the caller then hands it to the device)."""
        key = (task_id, parent_id)
        if self._vector is not None and self._valid_for == key:
            return False
        pattern = self.FILE.format(t=task_id, p=parent_id, i='[0-9]*?')
        while True:
            found = find_file_with_pattern(pattern, directory)
            if not found:
                break
            try:
                self._vector, self._valid_for = torch.load(os.path.join(directory, found), map_location='cpu'), key
                logging.info('took sensitivity from file')
                return True
            except (RuntimeError, EOFError):            # another worker is still writing it: wait and look again (:50-52)
                time.sleep(2)
        t0 = time.time()
        self._orig_batch_size = self._orig_batch_size or batch_size
        vec = self._calc_sensitivity(experiences).cpu().clamp_(min=self._underflow)
        vec /= self._underflow                           # :65 -- by the underflow value here, not by the minimum as in set_sensitivity
        logging.info('Safe mutation sensitivity computed in {:.2f}s on {} samples'.format(time.time() - t0, batch_size))
        if not find_file_with_pattern(pattern, directory):
            torch.save(vec.clone(), os.path.join(directory, self.FILE.format(t=task_id, p=parent_id, i=os.getpid())))
        self._vector, self._valid_for = vec, key
        logging.info('Sensitivity parent {}: min {:.2f}, mean {:.2f}, max {:.2f}'.format(parent_id, vec.min().item(), vec.mean().item(), vec.max().item()))
        return True

    def _calc_sensitivity(self, experiences):
        net = self.net
        theta = net.parameter_vector().to(device=net.ctx.dev, dtype=torch.float32)
        fwd = net.sensitivity_forward(experiences, self._orig_batch_size)
        return (sum_sensitivity_of if self._type.value == 'SM-G-SUM' else abs_sensitivity_of)(theta, fwd)
