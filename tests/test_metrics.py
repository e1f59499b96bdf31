"""Row N4: the string metrics of evaluate() (cor_asv_ann_b200/metrics.py) against the reference's own
lib/alignment.py, executed live from /root/reference where it exists (build container) and against the outputs
frozen from it in tests/golden/ref_metrics.json everywhere else; plus gen_lines against the reference's generator
and evaluate() end to end on the host (engine replaced by a deterministic fake corrector).
"""
import importlib.util
import json
import os
import pickle
import random
import sys
import types

import numpy as np
import pytest

from cor_asv_ann_b200 import metrics
from tests.common import GOLDEN_DIR, kant_fixture

REF_ALIGNMENT = "/root/reference/ocrd_cor_asv_ann/lib/alignment.py"
FIXTURE = os.path.join(GOLDEN_DIR, "ref_metrics.json")


def fresh_reference():
    """A FRESH import of the reference's alignment module (its normalisation table is consumed by the first call,
    so every scenario needs its own copy)."""
    if "uniseg" not in sys.modules:
        wb = types.ModuleType("uniseg.wordbreak")
        pkg = types.ModuleType("uniseg")
        pkg.wordbreak = wb
        sys.modules["uniseg"], sys.modules["uniseg.wordbreak"] = pkg, wb
    spec = importlib.util.spec_from_file_location("cor_ref_alignment_fresh", REF_ALIGNMENT)
    mod = importlib.util.module_from_spec(spec)
    sys.dont_write_bytecode = True
    spec.loader.exec_module(mod)
    return mod


def line_pairs():
    """(OCR-like, ground truth) pairs: the kant lines with seeded edits, historic ligatures, combining marks."""
    _, pages = kant_fixture()
    rng = random.Random(5)
    truth = [''.join(chunk[0][0] for chunk in line).replace('\a', '') for page in pages for line in page][:40]
    extra = ["Aufklaͤrung iſt der Ausgang\n", "daﬅ iﬁ  qū āb\n", "éä oͤ̈x\n",
             "„Zitat“ — 1² 3⁄ 4\n", "\n", "a\n"]
    pairs = []
    for t in truth + extra:
        chars = list(t[:-1])
        for _ in range(rng.randint(0, 4)):
            if not chars:
                break
            k = rng.randrange(len(chars))
            op = rng.random()
            if op < 0.4:
                chars[k] = rng.choice("aeinsſͤ̈rt .")
            elif op < 0.7:
                del chars[k]
            else:
                chars.insert(k, rng.choice("aeioͤ m"))
        pairs.append((''.join(chars) + '\n', t))
    return pairs


SCENARIOS = [("historic_latin", 1), ("historic_latin", 2), ("historic_latin", 3), ("NFC", 1), ("NFKC", 1), (None, 1)]


def run_ours(normalization, gt_level):
    rules = metrics.HistoricLatinRules()
    metrics._RULES = rules
    c_al, w_al = metrics.Alignment(0, confusion=True), metrics.Alignment(0)
    c_ed, w_ed = metrics.Edits(histogram=True), metrics.Edits()
    rows = []
    for src, tgt in line_pairs():
        d, n, al = c_al.get_adjusted_distance(src, tgt, normalization=normalization, gtlevel=gt_level, return_alignment=True)
        c_ed.add(d, n, src, tgt)
        ws, wt = metrics.splitwords(src), metrics.splitwords(tgt)
        dw, nw = w_al.get_adjusted_distance(ws, wt, normalization=normalization, gtlevel=gt_level)
        w_ed.add(dw, nw, ws, wt)
        rows.append([d, n, [[a or 0, b or 0] for a, b in al], dw, nw])
    conf, total = c_al.get_confusion(10)
    return dict(rows=rows, confusion=[[c, list(p)] for c, p in conf], total=total,
                c=[c_ed.length, c_ed.mean, c_ed.varia, c_ed.steps], w=[w_ed.length, w_ed.mean, w_ed.varia, w_ed.steps],
                hist=[[k, list(v)] for k, v in c_ed.hist().items()],
                worst=[[r, n] for _, r, n, _ in c_ed.worst])


def run_reference(normalization, gt_level):
    ref = fresh_reference()
    c_al, w_al = ref.Alignment(0, confusion=True), ref.Alignment(0)
    c_ed, w_ed = ref.Edits(histogram=True), ref.Edits()
    rows = []
    for src, tgt in line_pairs():
        d, n, al = c_al.get_adjusted_distance(src, tgt, normalization=normalization, gtlevel=gt_level, return_alignment=True)
        c_ed.add(d, n, src, tgt)
        ws, wt = metrics.splitwords(src), metrics.splitwords(tgt)      # uniseg is absent: same tokens for both sides
        dw, nw = w_al.get_adjusted_distance(ws, wt, normalization=normalization, gtlevel=gt_level)
        w_ed.add(dw, nw, ws, wt)
        rows.append([d, n, [[a or 0, b or 0] for a, b in al], dw, nw])
    conf, total = c_al.get_confusion(10)
    return dict(rows=rows, confusion=[[c.count, list(c.pair)] for c in conf], total=total,
                c=[c_ed.length, c_ed.mean, c_ed.varia, c_ed.steps], w=[w_ed.length, w_ed.mean, w_ed.varia, w_ed.steps],
                hist=[[k, list(v)] for k, v in c_ed.hist().items()],
                worst=[[e.mean, e.length] for e in c_ed.worst])


def _key(normalization, gt_level):
    return "%s/%d" % (normalization, gt_level)


@pytest.mark.parametrize("normalization,gt_level", SCENARIOS)
def test_metrics_match_reference_fixture(normalization, gt_level):
    want = json.load(open(FIXTURE, encoding="utf-8"))[_key(normalization, gt_level)]
    got = json.loads(json.dumps(run_ours(normalization, gt_level)))
    assert got == want


@pytest.mark.skipif(not os.path.isfile(REF_ALIGNMENT), reason="needs /root/reference")
@pytest.mark.parametrize("normalization,gt_level", SCENARIOS)
def test_metrics_match_reference_live(normalization, gt_level):
    assert json.loads(json.dumps(run_ours(normalization, gt_level))) == json.loads(json.dumps(run_reference(normalization, gt_level)))


def test_normalisation_table_is_consumed_by_the_first_string():
    """The reference pops the single-codepoint rules on first use (alignment.py:316-320)."""
    rules = metrics.HistoricLatinRules()
    assert metrics.normalize("ﬁ", "historic_latin", 1, rules) == "fi"
    assert metrics.normalize("ﬁ", "historic_latin", 1, rules) == "ﬁ"
    assert metrics.normalize("q̈", "historic_latin", 1, rules) == "qᷓ"       # multi-codepoint rules stay


@pytest.mark.skipif(not os.path.isfile(REF_ALIGNMENT), reason="needs /root/reference")
def test_equivalence_tables_are_the_reference_tables():
    ref = fresh_reference()
    assert [frozenset(c.split("|")) for c in metrics._L1_CLASSES] == [frozenset(s) for s in ref.L1_HISTLAT_EQV]
    assert list(metrics._L2_RULES) == list(ref.L2_HISTLAT_EQV.items())


def test_uax29_words():
    w = metrics.splitwords
    assert w("Was ist Aufklärung?\n") == ["Was", "ist", "Aufklärung"]
    assert w("can't 3,5 1.000,5 a.b a:b foo_bar x-y") == ["can't", "3,5", "1.000,5", "a.b", "a:b", "foo_bar", "x", "y"]
    assert w("été aͤb  — »«") == ["été", "aͤb"]
    assert w("z. B. 12te") == ["z", "B", "12te"]
    assert list(metrics.uax29_words("a  b\r\nc")) == ["a", "  ", "b", "\r\n", "c"]


# ----------------------------------------------------------------------------- gen_lines against the reference

def _write_inputs(tmp_path):
    rng = random.Random(3)
    text = tmp_path / "a.txt"
    rows = []
    for i in range(23):
        src = ''.join(rng.choice("abcde ſü") for _ in range(rng.randint(1, 12)))
        rows.append(src + "\t" + src.upper() + "\n")
    text.write_text(''.join(rows), encoding="utf-8")
    plain = tmp_path / "b.txt"
    plain.write_text("eins\nzwei\tZWEI\ndrei\n", encoding="utf-8")
    pk = tmp_path / "c.pkl"
    items = [([('a', 0.9), ('b', 0.5), ('\n', 1.0)], "ab\n"), ("xyz\n", "xyz\n"), ("", "leer\n"),
             ([[('a', 0.7), ('bc', 0.2)], [], [('\n', 1.0)]], "a\n"), ([('q', 0.4)], "q\n")]
    with open(pk, "wb") as f:
        pickle.dump(items, f)
    return str(text), str(plain), str(pk)


@pytest.mark.skipif(not os.path.isfile(REF_ALIGNMENT), reason="needs /root/reference")
def test_gen_lines_matches_reference_generator(tmp_path):
    from tests import ref_harness as rh
    from cor_asv_ann_b200.seq2seq import Sequence2Sequence
    ref_cls = rh.load_lib().Sequence2Sequence
    text, plain, pk = _write_inputs(tmp_path)
    split = np.random.default_rng(1).uniform(0, 1, 64)
    cases = [dict(filenames=[text], repeat=False), dict(filenames=[text, text], repeat=False, charmap={'a': 'A', 'ſ': 's'}),
             dict(filenames=[plain], repeat=False, unsupervised=True), dict(filenames=[pk], repeat=False),
             dict(filenames=[pk, pk], repeat=False, unsupervised=True),
             dict(filenames=[text], repeat=False, split=split, train=False), dict(filenames=[text], repeat=True, split=split, train=True)]
    for kw in cases:
        ours, ref = Sequence2Sequence(), ref_cls.__new__(ref_cls)
        ref.batch_size = ours.batch_size = 4
        ref.logger, ref.progbars = ours.logger, False
        ref.aligner = rh._MODS["alignment"].Alignment(0)
        g1, g2 = ours.gen_lines(**kw), ref.gen_lines(**kw)
        for _ in range(30 if kw["repeat"] else 10 ** 6):
            a, b = next(g1, "end"), next(g2, "end")
            assert a == b, kw
            if a == "end":
                break


# ----------------------------------------------------------------------------- evaluate() on the host

class _FakeCorrector:
    """correct_batches stand-in: a deterministic 'model' that fixes some characters and breaks others."""

    def __call__(self, batches, fast=True, greedy=True, want_align=False, want_probs=True):
        out = []
        for lines, _ in batches:
            fixed = [(l.replace('ſ', 's') if greedy else l.replace('ü', 'ü')) for l in lines]
            out.append((fixed, [[] for _ in lines], [0.1 * len(l) for l in lines], [[] for _ in lines]))
        return out


def _reference_evaluate(ref, batches, corrector, normalization, gt_level):
    """The reference's evaluate loop (seq2seq.py:684-734) on its own Alignment/Edits, fed the same predictions."""
    names = ['c_origin', 'c_greedy', 'c_beamed', 'w_origin', 'w_greedy', 'w_beamed']
    counts = {n: ref.Edits(histogram=n[0] == 'c') for n in names}
    al = {n: ref.Alignment(0, confusion=n[0] == 'c') for n in names}
    for batch in batches:
        src, conf, tgt, _ = batch
        g = corrector([(src, conf)], greedy=True)[0]
        m = corrector([(src, conf)], greedy=False)[0]
        for j in range(len(src)):
            if not src[j] or not tgt[j]:
                continue
            texts = dict(origin=src[j], greedy=g[0][j], beamed=m[0][j])
            for k in ('origin', 'greedy', 'beamed'):
                d, n = al['c_' + k].get_adjusted_distance(texts[k], tgt[j], normalization=normalization, gtlevel=gt_level)
                counts['c_' + k].add(d, n, texts[k], tgt[j])
            toks = {k: metrics.splitwords(v) for k, v in texts.items()}
            tt = metrics.splitwords(tgt[j])
            for k in ('origin', 'greedy', 'beamed'):
                d, n = al['w_' + k].get_adjusted_distance(toks[k], tt, normalization=normalization, gtlevel=gt_level)
                counts['w_' + k].add(d, n, toks[k], tt)
        counts['c_greedy'].score += sum(g[2])
        counts['c_beamed'].score += sum(m[2])
    return counts, al


@pytest.mark.skipif(not os.path.isfile(REF_ALIGNMENT), reason="needs /root/reference")
@pytest.mark.parametrize("workers", [0, 3])
def test_evaluate_equals_reference_loop(tmp_path, workers):
    from cor_asv_ann_b200.seq2seq import Sequence2Sequence
    text, _, _ = _write_inputs(tmp_path)
    big = tmp_path / "big.txt"
    big.write_text(open(text, encoding="utf-8").read() * 3 + "daﬅ iﬁ\tdaſt ifi\n", encoding="utf-8")
    s = Sequence2Sequence()
    s.batch_size, s.superbatch, s.prefetch, s.metric_workers, s.status = 8, 2, False, workers, 2
    s.correct_batches = _FakeCorrector()
    metrics._RULES = metrics.HistoricLatinRules()
    s.evaluate([str(big)], fast=False, normalization='historic_latin', gt_level=1, confusion=5, histogram=True)
    ref = fresh_reference()
    counts, al = _reference_evaluate(ref, list(s.gen_lines([str(big)], False)), _FakeCorrector(), 'historic_latin', 1)
    for name, c in counts.items():
        got = s.last_evaluation[name]
        assert (got["length"], got["steps"]) == (c.length, c.steps)
        assert got["mean"] == c.mean and got["varia"] == c.varia and got["score"] == c.score, name
    for k in ('origin', 'greedy', 'beamed'):
        table, total = al['c_' + k].get_confusion(5)
        assert s.last_evaluation['confusion'][k] == ([(c.count, c.pair) for c in table], total)


if __name__ == "__main__":      # regenerate the fixture from the reference (build container only)
    out = {_key(n, g): run_reference(n, g) for n, g in SCENARIOS}
    with open(FIXTURE, "w", encoding="utf-8") as f:
        json.dump(out, f, ensure_ascii=True)
    print("wrote", FIXTURE, os.path.getsize(FIXTURE), "bytes")
