"""``Sequence2Sequence.train`` (ocrd_cor_asv_ann/lib/seq2seq.py:592-650) on the CUDA trainer: the epoch loop of
``fit_generator_autosized`` (lib/keras_train.py) with the callbacks the reference installs -- early stopping on the
validation loss (patience 3, best weights restored), a weights checkpoint per epoch, stop on NaN -- over the batches of
``gen_lines`` (same files, same split, same batches).  Under ``torch.distributed`` every rank takes a contiguous share
of the rows of EVERY batch (vectorised to the batch's common padded lengths first: the model does not mask padding, so
the lengths are part of the input) and the gradients are all-reduced: the update equals the single-process one.

Not carried over: scheduled sampling (decoder-sampled teacher input without gradient, :860-905) and the LM loss.
"""
import unicodedata
import pickle

import numpy as np

from .vectorize import lines_to_sparse

GAP = '\a'


def map_files(s2s, filenames):
    """seq2seq.py:554-590: collect the characters of the files, extend the mapping (incremental training keeps the old
    indexes only if the sorted order allows -- the reference rebuilds the mapping from the sorted set), count lines."""
    num_lines = 0
    chars = set(s2s.mapping[0].keys())          # includes '' (0)
    for filename in filenames:
        with_confidence = filename.endswith('.pkl')
        with open(filename, 'rb' if with_confidence else 'r') as handle:
            items = pickle.load(handle) if with_confidence else handle
            for line in items:
                if with_confidence:
                    source_conf, target_text = line
                    if not source_conf:
                        line = target_text
                    elif type(source_conf[0]) is tuple:
                        line = ''.join(char for char, prob in source_conf) + target_text
                    else:
                        line = ''.join(chars_ for chunk in source_conf for chars_, prob in chunk) + target_text
                line = unicodedata.normalize('NFC', line)
                chars.update(set(line))
                if GAP in chars:
                    s2s.logger.warning('ignoring gap character "%s" in input file "%s"', GAP, filename)
                    chars.remove(GAP)
                num_lines += 1
    chars = sorted(chars)
    if len(chars) > s2s.voc_size:
        c_i = {c: i for i, c in enumerate(chars)}
        i_c = {i: c for i, c in enumerate(chars)}
        s2s.mapping = (c_i, i_c)
        s2s.voc_size = len(c_i)
        s2s._reconfigure_for_mapping()
    return num_lines


def batch_arrays(s2s, sources, confs, targets, rng=None, degrade=False):
    """vectorize_lines (seq2seq.py:1020-1117) as index arrays + the training-time input degradation (:909-915):
    (src_idx [B,Ts,k], src_val or None, dec_in [B,Td], dec_out [B,Td])."""
    c2i = s2s.mapping[0]
    idx, val = lines_to_sparse(c2i, sources, confs, s2s.logger)
    B, Ts, k = idx.shape
    Td = max(map(len, targets)) + 1
    di = np.full((B, Td), -1, np.int32)
    do = np.full((B, Td), -1, np.int32)
    for i, t in enumerate(targets):
        ids = [c2i.get(ch, 0) for ch in t]
        di[i, 1:len(t) + 1] = ids
        do[i, :len(t)] = ids
    if degrade and rng is not None:
        pos = (Ts * rng.uniform(0, 1, B) / 0.01).astype(np.int64)
        rows = np.nonzero(pos < Ts)[0]
        if len(rows):
            idx[rows, pos[rows], :] = -1
            idx[rows, pos[rows], 0] = 0              # the index-0 unit vector
            if val is not None:
                val[rows, pos[rows], 0] = 1.0
    return idx, val, di, do


def _share(n_rows, rank, world):
    lo = (n_rows * rank) // world
    return lo, (n_rows * (rank + 1)) // world


def run_epoch(s2s, trainer, batches, rank, world, train, rng):
    """One pass; returns (mean loss over batches weighted by batch size like Keras, batches)."""
    total, count, n_batches = 0.0, 0, 0
    for batch in batches:
        if batch is False:
            break
        sources, confs, targets, _ = batch
        idx, val, di, do = batch_arrays(s2s, sources, confs, targets, rng, degrade=train)
        lo, hi = _share(len(sources), rank, world)
        sl = slice(lo, hi)
        masks = trainer.dropout_masks(rng, s2s.dropout, hi - lo) if train else None
        if hi > lo:
            xent, n = trainer.forward_backward(idx[sl], None if val is None else val[sl], di[sl], do[sl], masks)
        else:
            trainer.grads.zero_()
            xent, n = 0.0, 0.0
        xent, n = trainer.all_reduce(xent, n)
        reg = 0.0
        if train:
            reg, _ = trainer.apply(n)
        loss = xent / max(n, 1.0) + reg
        if loss != loss:
            return float('nan'), n_batches      # TerminateOnNaN
        total += loss * len(sources)
        count += len(sources)
        n_batches += 1
    return (total / count if count else float('nan')), n_batches


def train(s2s, filenames, val_filenames=None, seed=None):
    import torch
    import torch.distributed as dist
    from .trainer import Trainer
    if s2s.scheduled_sampling:
        raise NotImplementedError('scheduled sampling is not part of the B200 training step')
    if s2s.lm_loss:
        raise NotImplementedError('the LM loss is not part of the B200 training step')
    rank, world = (dist.get_rank(), dist.get_world_size()) if (dist.is_available() and dist.is_initialized()) else (0, 1)
    num_lines = s2s.map_files(filenames)
    s2s.logger.info('Training on "%d" files with %d lines', len(filenames), num_lines)
    rng = np.random.default_rng(seed)
    if val_filenames:
        num_lines = s2s.map_files(val_filenames)
        s2s.logger.info('Validating on "%d" files with %d lines', len(val_filenames), num_lines)
        split_rand = None
    else:
        s2s.logger.info('Validating on random 20% lines from those files')
        split_rand = rng.uniform(0, 1, (num_lines,))
        if world > 1:                       # every rank must reserve the same lines
            t = torch.as_tensor(split_rand, device=torch.device("cuda", s2s.device))
            dist.broadcast(t, 0)
            split_rand = t.cpu().numpy()
    if s2s.status < 1:
        s2s.configure()
    topo = dict(width=int(s2s.width), depth=int(s2s.depth), voc_size=int(s2s.voc_size),
                residual_connections=bool(s2s.residual_connections), deep_bidirectional_encoder=bool(s2s.deep_bidirectional_encoder),
                bridge_dense=bool(s2s.bridge_dense), window_width=5)
    trainer = Trainer(topo, s2s._weights, device=s2s.device)
    if world > 1:                           # identical start everywhere
        dist.broadcast(trainer.params, 0)
    best, best_params, wait, history = float('inf'), None, 0, []
    stopped_epoch = None
    for epoch in range(int(s2s.epochs)):
        loss, n_train = run_epoch(s2s, trainer, s2s.gen_lines(filenames, True, split_rand, True), rank, world, True, rng)
        val_loss, _ = run_epoch(s2s, trainer, s2s.gen_lines(val_filenames or filenames, True, split_rand, False), rank, world, False, rng)
        history.append(val_loss)
        s2s.logger.info('epoch %d: %d batches, loss %.4f, val_loss %.4f', epoch + 1, n_train, loss, val_loss)
        if loss != loss or val_loss != val_loss:
            s2s.logger.warning('invalid loss, terminating training')
            break
        if rank == 0:
            s2s._weights = trainer.get_weights()
            s2s._engine = None
            s2s._save_weights("model.ckpt.weights-%02d-%.2f.h5" % (epoch + 1, val_loss))
        if val_loss < best:
            best, best_params, wait = val_loss, trainer.params.clone(), 0
        else:
            wait += 1
            if wait >= 3:                   # EarlyStopping(patience=3, restore_best_weights=True)
                stopped_epoch = epoch
                break
    if history and best_params is not None:
        s2s.logger.info('training finished with val_loss %f', min(history))
        trainer.params.copy_(best_params)
        s2s._weights = trainer.get_weights()
        s2s._engine = None
        s2s._resync_decoder()
        s2s.status = 2
    else:
        s2s.logger.critical('training failed')
        s2s.status = 1
    s2s.history = {"val_loss": history, "stopped_epoch": stopped_epoch}
    trainer.close()
