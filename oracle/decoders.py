"""Oracle, part 2: input vectorisation and the three decoders (numpy + host Python).

TEST INFRASTRUCTURE — see ``oracle/__init__.py``.  Parity unpinned.

Restates (file:line under /root/reference/ocrd_cor_asv_ann/lib/seq2seq.py):
  * ``vectorize_lines``          :1020-1119  (encoder side only)
  * ``decode_batch_greedy``      :1215-1286
  * ``decode_sequence_greedy``   :1288-1354
  * ``decode_sequence_beam``     :1356-1544  and ``Node`` :1546-1608
  * ``correct_lines``            :782-842

Numerical-type notes that matter for bit-parity of the beam order:
the reference ran under numpy<1.24, where ``python_float + np.float32`` and
``np.float32 * python_float`` promote to float64.  numpy>=2 (this image) keeps
float32, so every such place casts explicitly here.
"""
import bisect
import logging
import numpy as np

from .model import encode, decode_step

GAP = '\a'
LOG = logging.getLogger("oracle")

# Scalar promotion rules the bookkeeping below follows.  "legacy" = numpy<1.24, the reference's
# pinned environment (Dockerfile:38): ``python_float + np.float32`` -> float64.  "nep50" = numpy>=2
# (this image): python scalars are weak, the result stays float32.  The switch exists so that
# tests/ref_harness.py can run the reference's own decoder code under this image's numpy and
# compare it with this restatement bit for bit; the product follows "legacy".
PROMOTION = "legacy"


def _nep50():
    return PROMOTION == "nep50"


# --------------------------------------------------------------------------- vectorisation

def vectorize_lines(mapping, voc_size, lines, conf=None):
    """Encoder half of seq2seq.py:1020-1119.

    ``lines``: list of str; ``conf``: None | list of per-char probability lists |
    list of confusion networks (line = list of chunks, chunk = list of
    (string, prob) alternatives).  Returns the dense (B,T,V) array —
    uint32 without ``conf``, float32 with.
    """
    c2i = mapping[0]
    confmat = bool(conf) and type(conf[0][0]) is list
    if confmat:
        seqs = conf
        T = max(sum(max(len(alt[0]) for alt in chunk) if chunk else 0 for chunk in line) for line in conf)
    else:
        seqs = lines
        T = max(len(s) for s in lines)
    x = np.zeros((len(seqs), T, voc_size), np.float32 if conf else np.uint32)
    for i, seq in enumerate(seqs):
        if confmat:
            j = 0
            for chunk in seq:
                width = max(len(alt[0]) for alt in chunk) if chunk else 0
                for chars, p in chunk:
                    for k, ch in enumerate(chars):
                        if ch not in c2i and ch != GAP:
                            LOG.error('unmapped character "%s" at encoder input sequence %d position %d', ch, i, j + k)
                        x[i, j + k, c2i.get(ch, 0)] = p
                j += width
        else:
            for j, ch in enumerate(seq):
                if ch not in c2i and ch != GAP:
                    LOG.error('unmapped character "%s" at encoder input sequence %d', ch, i)
                x[i, j, c2i.get(ch, 0)] = conf[i][j] if conf else 1
    return x


# --------------------------------------------------------------------------- fast / batch greedy

def decode_batch_greedy(topo, w, mapping, x_in, hoist_u=True, max_steps=None, trace=None):
    """seq2seq.py:1215-1286.  Always 2T decoder steps; argmax over v>=1; softmax feedback.

    Returns (strings, probs, scores, alignments) — the reference's first return
    value (the uint32 copy of the fed-back inputs, used only by scheduled
    sampling) is not reproduced.
    ``hoist_u=False`` recomputes enc_out·U inside every step like the reference
    (used for the CPU-baseline timing); results are identical.
    ``trace``: optional list that receives the (B,V) score matrix of every step.
    """
    i2c = mapping[1]
    enc = encode(topo, w, x_in)
    enc_out, states = enc[0], enc[1:]
    B, T = x_in.shape[0], x_in.shape[1]
    u = enc_out @ w["attention_dense"][0] if hoist_u else None
    p = np.zeros((B, topo.voc_size), enc_out.dtype)
    live_input = np.asarray(x_in).reshape(B, -1).any(axis=1)
    text = [''] * B
    probs = [[] for _ in range(B)]
    score = [0.] * B
    aligns = [[] for _ in range(B)]
    for _ in range(2 * T if max_steps is None else max_steps):
        p, _, states = decode_step(topo, w, p, enc_out, states, u=u)
        if trace is not None:
            trace.append(p.copy())
        best = np.nanargmax(p[:, 1:], axis=1) + 1
        with np.errstate(divide='ignore'):
            nlog = -np.log(p)
        a = states[-1]
        for j in range(B):
            if text[j].endswith('\n') or not live_input[j]:
                continue
            v = int(best[j])
            text[j] += i2c[v]
            probs[j].append(p[j, v])
            if _nep50():
                score[j] = np.float32(score[j]) + nlog[j, v]
            else:
                score[j] += float(nlog[j, v])     # python float + np.float32 -> float64 under numpy<1.24
            aligns[j].append(a[j])
    for j in range(B):
        if text[j]:
            score[j] /= len(text[j])
    return text, probs, score, aligns


# --------------------------------------------------------------------------- per-line greedy

def decode_sequence_greedy(topo, w, mapping, enc_slices, events=None):
    """seq2seq.py:1288-1354 with B=1; argmax over ALL v, re-argmax if '' wins.

    Deviation: the reference feeds the NaN-patched vector
    back, which makes the next step raise; here the original softmax row is fed
    back and the event is recorded in ``events``.
    """
    i2c = mapping[1]
    enc_out, states = enc_slices[0], list(enc_slices[1:])
    T = enc_out.shape[1]
    u = enc_out @ w["attention_dense"][0]
    p = np.zeros((1, topo.voc_size), enc_out.dtype)
    # ``decoded_score = 0; decoded_score += -np.log(prob)``: int + np.float32 is float64 under the reference's numpy
    # (< 1.24, legacy promotion), float32 under numpy >= 2 (PROMOTION "nep50", used by tests/ref_harness.py)
    acc = np.float32 if _nep50() else np.float64
    text, probs, total, aligns = '', [], acc(0), []
    for step in range(2 * T):
        p, _, states = decode_step(topo, w, p, enc_out, states, u=u)
        row = p[0]
        v = int(np.nanargmax(row))
        if i2c[v] == '':
            if events is not None:
                events.append(('nul-argmax', step))
            patched = row.copy()
            patched[v] = np.nan
            v = int(np.nanargmax(patched))
        text += i2c[v]
        probs.append(row[v])
        total = acc(total + acc(-np.log(row[v])))
        aligns.append(states[-1][0])
        if i2c[v] == '\n':
            break
    return text, probs, total / len(text), aligns


# --------------------------------------------------------------------------- beam (A*)

class Node:
    """Trie node of the beam (seq2seq.py:1546-1608)."""
    __slots__ = ("value", "parent", "state", "cum_cost", "length", "length0", "cost0",
                 "prob", "scores", "alignment", "serial")

    def __init__(self, state, value, scores, cost, parent=None, prob=1.0, alignment=None,
                 length0=None, cost0=None):
        self.value = value
        self.parent = parent
        self.state = state
        # float64 accumulation of the float32 -log p (legacy numpy promotion)
        if _nep50():
            self.cum_cost = np.float32(parent.cum_cost) + np.float32(cost) if parent else cost
        else:
            self.cum_cost = np.float64(parent.cum_cost) + np.float64(cost) if parent else np.float64(cost)
        self.length = 1 if parent is None else parent.length + 1
        self.length0 = length0 or (parent.length0 if parent else 1)
        self.cost0 = cost0 or (parent.cost0 if parent else 0)
        self.prob = prob
        self.scores = scores
        self.alignment = (parent.alignment if parent else []) if alignment is None else alignment
        self.serial = 0

    def pro_cost(self):
        return -(self.cum_cost + self.cost0 * np.abs(self.length - self.length0))

    def path(self):
        out, n = [], self
        while n is not None:
            out.append(n)
            n = n.parent
        return out[::-1]


class _Fringe:
    """Ascending list by pro_cost; a new node goes LEFT of equal ones and the best is
    popped from the right (``insort_left`` + ``pop()``), i.e. FIFO among equal keys."""

    def __init__(self):
        self.keys, self.nodes, self.count = [], [], 0

    def push(self, node):
        self.count += 1
        node.serial = self.count
        k = (float(node.pro_cost()), -self.count)
        pos = bisect.bisect_left(self.keys, k)
        self.keys.insert(pos, k)
        self.nodes.insert(pos, node)

    def pop(self):
        self.keys.pop()
        return self.nodes.pop()

    def __len__(self):
        return len(self.nodes)

    def keep_best(self, n):
        if len(self.nodes) > n:
            self.keys = self.keys[-n:]
            self.nodes = self.nodes[-n:]


def decode_sequence_beam(topo, w, mapping, source_seq, enc_slices, *, batch_size=256,
                         beam_width_in=15, beam_threshold_in=0.2, beam_width_out=16,
                         rejection_threshold=0.3, stats=None):
    """seq2seq.py:1356-1544.  Generator of (string, probs, cum_cost/(len-1), alignments), best first.

    ``source_seq`` (T,V) dense input row block of this line (padded to the batch T),
    ``enc_slices`` = that line's [j:j+1] slices of ``encode(...)``.
    ``stats`` (dict) receives iterations / hypothesis-steps / tie diagnostics.
    """
    i2c = mapping[1]
    V = topo.voc_size
    enc_out, states0 = enc_slices[0], list(enc_slices[1:])
    T = enc_out.shape[1]
    u = enc_out @ w["attention_dense"][0]
    positions = np.arange(T)
    fringe, finals = _Fringe(), _Fringe()
    fringe.push(Node(state=states0, value='', scores=np.zeros(V), prob=[], cost=0.0,
                     alignment=[], length0=T, cost0=3.0))
    max_batches = 2 * T
    n_iter = n_hyp = n_score_ties = n_near = 0
    pop_sum, fringe_peak = 0, 1
    c2i = mapping[0]
    for _ in range(max_batches):
        beam = []
        while len(fringe):
            node = fringe.pop()
            if node.value == '\n':
                finals.push(node)
            else:
                beam.append(node)
                if node.length > 1.5 * T:
                    LOG.warning('found overlong hypothesis')
            if len(beam) >= batch_size:
                break
        if not beam:
            break
        if len(finals) > beam_width_out and finals.nodes[-1].pro_cost() > beam[0].pro_cost():
            break
        if stats is not None:
            # diagnostics the CUDA search reports as stats[5..7] (include/cor_b200.h): pop-boundary near-ties, a
            # checksum of which hypotheses are expanded in which iteration, the peak fringe length
            for t, nd in enumerate(beam):
                xx = (int(nd.length) << 16) + (c2i[nd.value] if nd.length > 1 else 0) + 1
                pop_sum = (pop_sum + xx * (2 * n_iter + 1) * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
            if len(fringe):
                ka, kb = -fringe.keys[-1][0], -float(beam[-1].pro_cost())
                if abs(ka - kb) <= 1e-6 * max(1.0, abs(ka)):
                    n_near += 1
        n_iter += 1
        n_hyp += len(beam)
        p_in = np.vstack([n.scores for n in beam]).astype(enc_out.dtype)
        st_in = [np.vstack([n.state[k] for n in beam]) for k in range(len(states0))]
        p_out, lm_out, st_out = decode_step(topo, w, p_in, enc_out, st_in, u=u)
        for i, node in enumerate(beam):
            st = [layer[i:i + 1] for layer in st_out]
            sc = p_out[i]                      # mutated in place below, like the reference
            al = st[-1][0]
            mis = 0.0
            if node.length > 1:
                prev = np.matmul(node.alignment, positions)          # float64
                cur = np.matmul(al, positions)                       # float64 (float32 @ int64)
                mis = np.abs(cur - prev - 1)
                if np.max(node.alignment) == 1.0:
                    pos = int(prev) + 1
                else:
                    pos = int(cur.round())
            else:
                pos = 0
            src = source_seq[pos]              # IndexError if pos == T: unguarded in the reference
            rej = None
            if (rejection_threshold
                    and (mis < 0.1 or np.max(node.alignment) == 1.0)
                    and np.any(src)):
                rej = int(np.nanargmax(src))
                thr = sc.dtype.type(rejection_threshold) if _nep50() else rejection_threshold
                if np.float64(sc[rej]) < thr:
                    sc[rej] = rejection_threshold
            order = np.argsort(sc)
            if _nep50():
                cut = sc[order[-1]] * sc.dtype.type(beam_threshold_in)
                beampos = V - int(np.searchsorted(sc[order], cut))
            else:
                hi = np.float64(sc[order[-1]])
                beampos = V - int(np.searchsorted(sc[order].astype(np.float64), hi * beam_threshold_in))
            beampos = min(beampos, beam_width_in)
            if stats is not None and V > 1 and sc[order[-1]] == sc[order[-2]]:
                n_score_ties += 1
            if stats is not None:
                # near-tie ordering decisions of this expansion (same definition as csrc/beam.cu, margin 1e-6 relative, + 2e-6 absolute at the threshold cut):
                # two consecutive walked candidates, the last one taken against the first left out, the threshold cut
                so = sc[order[::-1]].astype(np.float64)
                bound = np.float64(so[0]) * beam_threshold_in
                nw = min(beampos + 1, V)
                tie = bool(np.any(np.abs(so - bound) <= 1e-6 * bound + 2e-6))
                tie = tie or bool(np.any(so[:nw - 1] - so[1:nw] <= 1e-6 * so[:nw - 1]))
                if node.length > 1:      # position rounding, misalignment threshold
                    onehot = np.max(node.alignment) == 1.0
                    tie = tie or (not onehot and abs(abs(cur - np.floor(cur)) - 0.5) <= 1e-5) or abs(mis - 0.1) <= 1e-6
                n_near += int(tie)
            rank = 0
            for v in order[::-1]:
                v = int(v)
                rank += 1
                prob = sc[v]
                with np.errstate(divide='ignore'):
                    cost = -np.log(lm_out[i][v]) if topo.lm_predict else -np.log(prob)
                al1 = al
                if v == rej:
                    al1 = np.eye(T)[pos]
                    rej = None
                elif rank > beampos:
                    if rej:                    # NB: rej == 0 counts as "not pending" (A.6 item 7)
                        continue
                    break
                ch = i2c[v]
                if np.isnan(cost) or ch == '':
                    continue
                feedback = sc.copy()
                sc[v] = 0
                fringe.push(Node(parent=node, state=st, value=ch, scores=feedback,
                                 prob=prob, cost=cost, alignment=al1))
        fringe.keep_best(max_batches * batch_size)
        fringe_peak = max(fringe_peak, len(fringe))
    if stats is not None:
        stats.update(iterations=n_iter, hypothesis_steps=n_hyp, score_ties=n_score_ties,
                     finals=len(finals), open_left=len(fringe), near_ties=n_near,
                     pop_checksum=pop_sum, fringe_peak=fringe_peak)
    while len(finals):
        node = finals.pop()
        nodes = node.path()[1:]
        yield (''.join(n.value for n in nodes), [n.prob for n in nodes],
               node.cum_cost / (node.length - 1), [n.alignment for n in nodes])


# --------------------------------------------------------------------------- dispatcher

def correct_lines(topo, w, mapping, lines, conf=None, fast=True, greedy=True, beam_params=None, stats=None):
    """seq2seq.py:782-842."""
    assert not fast or greedy, "cannot decode in fast mode with beam search enabled"
    if not lines:
        return [], [], [], []
    x = vectorize_lines(mapping, topo.voc_size, lines, conf)
    if fast:
        return decode_batch_greedy(topo, w, mapping, x)
    enc = encode(topo, w, x)
    out = ([], [], [], [])
    for j, line in enumerate(lines):
        sl = [e[j:j + 1] for e in enc]
        if not line:
            res = ('', [], 0, [])
        elif greedy:
            res = decode_sequence_greedy(topo, w, mapping, sl)
        else:
            st = {} if stats is not None else None
            try:
                res = next(decode_sequence_beam(topo, w, mapping, x[j], sl, stats=st, **(beam_params or {})))
            except StopIteration:
                LOG.error('cannot beam-decode input line %d', j)
                if isinstance(line[0], tuple):
                    text = ''.join(chunk[0] for chunk in line)
                if isinstance(line[0], list):
                    text = ''.join(chunk[0][0] if chunk else '' for chunk in line)
                else:
                    text = line
                res = (text, [1.0] * len(text), 0, np.eye(len(text)).tolist())
            if stats is not None:
                stats.setdefault("lines", []).append(st)
        text = res[0].replace(GAP, '')
        for acc, item in zip(out, (text,) + tuple(res[1:])):
            acc.append(item)
    return out
