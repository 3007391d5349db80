"""Validation decode with the reference's interface (src/captioning/eval_utils.py:13-27,60-107): greedy captions for one
feature per image over a split of the reference's loader, handed to a language evaluator.

The reference's evaluator (language_eval, eval_utils.py:30-57) is pycocoevalcap: Java jars and the COCO annotation file, neither
of which exists here, so it is injected: language_eval(predictions, directory, split) -> {metric: score}.  The decode -- the
part that runs the model -- goes through the same device path as training rollouts."""
import numpy as np
import torch


def decode_sequence(ix_to_word, seq):
    """token ids -> words up to the first 0 (eval_utils.py:13-27); seq int [N, T] tensor or array"""
    seq = np.asarray(seq.cpu() if torch.is_tensor(seq) else seq)
    out = []
    for i in range(seq.shape[0]):
        words = []
        for j in range(seq.shape[1]):
            ix = int(seq[i, j])
            if ix <= 0:
                break
            words.append(ix_to_word[str(ix)])
        out.append(' '.join(words))
    return out


def eval_split(model, loader, directory, num=-1, split='val', do_eval=True, verbose=False, incl_gts=False, language_eval=None):
    """-> (lang_stats or None, predictions [{'image_id', 'caption'[, 'gts']}]) like the reference's eval_split.
    loader: the reference's DataLoader protocol (reset_iterator, get_batch, batch_size, seq_per_img, get_vocab)."""
    loader.reset_iterator(split)
    n = 0
    predictions = []
    while True:
        data = loader.get_batch(split)
        n = n + loader.batch_size
        # one feature per image: the loader replicates every image seq_per_img times (eval_utils.py:74-77)
        fc_feats = torch.from_numpy(np.ascontiguousarray(data['fc_feats'][np.arange(loader.batch_size) * loader.seq_per_img]))
        seq = model(fc_feats, greedy=True)[0]
        sents = decode_sequence(loader.get_vocab(), seq)
        for k, sent in enumerate(sents):
            entry = {'image_id': data['infos'][k]['id'], 'caption': sent}
            if incl_gts:
                entry['gts'] = decode_sequence(loader.get_vocab(), np.asarray(data['gts'][k]).astype(np.int32))
            predictions.append(entry)
        ix1 = data['bounds']['it_max']
        if num != -1:
            ix1 = min(ix1, num)
        for _ in range(n - ix1):
            predictions.pop()
        if data['bounds']['wrapped']:
            break
        if num >= 0 and n >= num:
            break
    lang_stats = None
    if do_eval:
        if language_eval is None:
            raise NotImplementedError('language_eval needs pycocoevalcap (Java) and the COCO annotations, which are not part of this '
                                      'build: pass language_eval=callable(predictions, directory, split) or do_eval=False')
        lang_stats = language_eval(predictions, directory, split)
    return lang_stats, predictions
