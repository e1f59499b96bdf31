"""String metrics behind ``Sequence2Sequence.evaluate``.

What the reference computes in ``ocrd_cor_asv_ann/lib/alignment.py`` (``Alignment`` :140-364, ``Edits`` :366-459,
``splitwords`` :461-485) and drives from ``lib/seq2seq.py:651-754``: per line, a difflib alignment of OCR / greedy /
beamed text against the ground truth on character and on word level, an edit count under one of three
normalisations, running length-weighted mean and variance of the rates, a confusion table and a character
histogram.  Once decoding runs at millions of characters per second this pure-Python loop is what
``cor-asv-ann-eval`` waits for, so it is organised for the host cores here:

* ``line_counts`` is a pure function of one line's four strings: it can run in worker processes
  (``LineScorer``), many lines at a time, while the GPU decodes the next batches;
* the results are folded into the running statistics in LINE ORDER by the caller, so every float (the running
  mean/variance recurrences are order dependent) and every table order equals the reference's serial loop.

Reference quirks reproduced on purpose (results must be identical):

* ``historic_latin`` normalisation: the reference pops the single-codepoint rules out of its module-level table
  on first use (``alignment.py:316-320``), so only the FIRST string normalised in a process sees them; every
  later string gets the multi-codepoint rules only.  ``HistoricLatinRules`` keeps that state explicitly
  (``single_pending``) instead of mutating a global.
* ``Edits.update`` divides the pooled variance by the new length after adding the length (:431-437).

Word segmentation: the reference calls ``uniseg.wordbreak.words`` (UAX #29).  ``uniseg`` is used when it is
importable; otherwise ``uax29_words`` below applies the UAX #29 word-boundary rules with the Word_Break
property derived from ``unicodedata`` categories plus the standard's explicit code point lists -- exact for
the Latin-script (historic German) material this tool is made for, approximate for scripts whose
Word_Break class is not a function of the general category.
"""
import bisect
import difflib
import logging
import unicodedata

# Equivalence data of the reference (alignment.py:8-46 and :48-137), same content in compact form:
# level-1 classes, members separated by "|"; level-2 rewrite rules in the reference's dict order (later
# duplicates of a key already folded in, as a Python dict literal does).
_L1_CLASSES = (
    "\xe4|a\u0308|a\u0364",
    "\xf6|o\u0308|o\u0364",
    "\xfc|u\u0308|u\u0364",
    "\xc4|A\u0308|A\u0364",
    "\xd6|O\u0308|O\u0364",
    "\xdc|U\u0308|U\u0364",
    "s|\u017f",
    "r|\ua75b",
    "z|\u0292",
    "Z|\u01b7",
    "n|\u019e",
    "\xb5|\u03bc",
    "\u03c0|\U0001d6d1|\U0001d70b|\U0001d745|\U0001d77f|\U0001d7b9",
    "0|\u2070",
    "1|\xb9",
    "2|\xb2",
    "3|\xb3",
    "4|\u2074",
    "5|\u2075",
    "6|\u2076",
    "7|\u2077",
    "8|\u2078",
    "9|\u2079|\ua770",
    "\xbb|\u201e|\u203a|\u301f",
    "\xab|\u201c|\u2039|\u301e",
    "'|`|\u02b9|\u02bc|\u1fbd|\u2018|\u2019|\u201b|\u2032",
    ",|\u201a",
    "-|\u2010|\u2011|\u2012|\u2013|\u2014|\u2015|\u2043|\u2212|\u2500|\u2e17|\ufe58",
    "\u201d|\u201f|\u2033|\u3003",
    "~|\u02dc|\u1fc0|\u2053|\u223c",
    "(|\u207d|\u27e8",
    ")|\u207e|\u27e9",
    "/|\u2044|\u2215|\u29f8",
    "\\|\u2216|\u29f5|\u29f9",
)
_L2_RULES = (
    ("\uf502", "ch"),
    ("\ueec4", "ck"),
    ("\ufb05", "\u017ft"),
    ("\ufb01", "fi"),
    ("\ufb00", "ff"),
    ("\ufb02", "fl"),
    ("\ufb03", "ffi"),
    ("\uf4fc", "\u017fk"),
    ("\ueedc", "t\u0292"),
    ("\uf532", "as"),
    ("\uf533", "is"),
    ("\uf534", "us"),
    ("\uf535", "Qu"),
    ("\u0133", "ij"),
    ("\ue8bf", "q\u0292"),
    ("\ueba5", "\u017fp"),
    ("\ufb06", "st"),
    ("q\u0308", "q\u1dd3"),
    ("c\u0308", "c\u1dd3"),
    ("\u1e21", "g\u1dd3"),
    ("v\u0309", "v\u1de3"),
    ("v\u1dce", "v\u1de3"),
    ("b\u1dce", "b\u1de3"),
    ("p\u1dce", "p\u1de3"),
    ("d\u0309", "\xf0"),
    ("\ua75f", "v\u1de3"),
    ("t\u1de3", "t\u1dd1"),
    ("\ueada", "\u017ft"),
    ("\ueba2", "\u017fi"),
    ("\ueba3", "\u017fl"),
    ("\ueba6", "\u017f\u017f"),
    ("\ueba7", "\u017f\u017fi"),
    ("\uf4ff", "\u017f\u017ft"),
    ("\uf52c", "\u017fp"),
    ("\ueec5", "ct"),
    ("\ueecb", "ft"),
    ("\ue5d2", "m\u0303"),
    ("\ue5dc", "\xf1"),
    ("\ue665", "p\u0303"),
    ("\ue42c", "a\u0364"),
    ("\ue644", "o\u0364"),
    ("\ue72b", "u\u0364"),
    ("\ue72d", "\u016f"),
    ("\uebac", "\xdf"),
    ("\ue8b7", "\u017f\u1de3"),
    ("\uf1a6", "\ua770"),
    ("\uf223", "m"),
    ("\uf158", "\u204a"),
    ("\uf159", "\xf0"),
    ("\uf160", ":"),
    ("q\uf02f", "q\u0365"),
    ("t\uf1cc", "t\u1dd1"),
    ("\uf4f9", "ll"),
    ("\u0101", "a\u0303"),
    ("\u0113", "\u1ebd"),
    ("\u012b", "\u0129"),
    ("\u014d", "\xf5"),
    ("\u016b", "\u0169"),
    ("c\u0304", "c\u0303"),
    ("q\u0304", "q\u0303"),
    ("r\u0304", "r\u0303"),
    ("\uf50e", "q\u0301"),
)


GAP_ELEMENT = 0


class HistoricLatinRules:
    """The two ``historic_latin`` levels: rewrite rules applied to each string before alignment (GT level < 3) and
    equivalence classes consulted when comparing aligned symbols (GT level 1)."""

    def __init__(self):
        self.classes = [frozenset(c.split("|")) for c in _L1_CLASSES]
        self.multi = [(k, v) for k, v in _L2_RULES if len(k) != 1]
        self.single = str.maketrans({k: v for k, v in _L2_RULES if len(k) == 1})
        self.single_pending = True          # see the module docstring: consumed by the first string normalised

    def rewrite(self, text):
        """One string through the level-2 rules; consumes the single-codepoint table on first use."""
        table = self.single if self.single_pending else None
        self.single_pending = False
        for k, v in self.multi:
            text = text.replace(k, v)
        return text.translate(table) if table else text

    def same_class(self, x, y):
        for c in self.classes:
            if x in c and y in c:
                return True
        return False


_RULES = HistoricLatinRules()       # per process, like the reference's module-level table


def normalize(seq, normalization, gt_level=1, rules=None):
    """``get_adjusted_distance``'s inner ``normalize`` (alignment.py:306-325): a string, or a list of tokens."""
    rules = rules or _RULES
    if isinstance(seq, list):
        return [normalize(s, normalization, gt_level, rules) for s in seq]
    if normalization in ('NFC', 'NFKC'):
        return unicodedata.normalize(normalization, seq)
    if normalization == 'historic_latin':
        if gt_level < 3:
            return rules.rewrite(seq)
        rules.single_pending = False      # the reference builds (and discards) an empty table here; state is untouched
        return seq
    return seq


def _equivalent(x, y, classes_on, rules):
    if isinstance(x, list):
        return len(x) == len(y) and all(_equivalent(a, b, classes_on, rules) for a, b in zip(x, y))
    if x == y:
        return True
    return classes_on and rules.same_class(x, y)


def align_pairs(source, target, gap=GAP_ELEMENT):
    """Pairs (source symbol | gap, target symbol | gap) from difflib's opcodes (Ratcliff-Obershelp blocks,
    ``autojunk`` off), alignment.py:164-201: a replace block pairs positionally and pads the longer side."""
    sm = difflib.SequenceMatcher(isjunk=None, autojunk=False)
    sm.set_seqs(source, target)
    pairs = []
    for tag, i1, i2, j1, j2 in sm.get_opcodes():
        n, m = i2 - i1, j2 - j1
        both = min(n, m) if tag in ('equal', 'replace') else 0
        pairs.extend(zip(source[i1:i1 + both], target[j1:j1 + both]))
        pairs.extend((s, gap) for s in source[i1 + both:i2])
        pairs.extend((gap, t) for t in target[j1 + both:j2])
    return pairs


def _is_letter_head(sym):
    return unicodedata.category(sym[0])[0] == 'L'


def recombine_graphemes(pairs, gap=GAP_ELEMENT):
    """alignment.py:203-241: a combining codepoint joins the previous position of its own side when that holds
    a letter; the partner symbol joins likewise, else it is left behind against a gap."""
    out, changed = [], False
    for s, t in pairs:
        if (s != gap and unicodedata.combining(s) and out and out[-1][0] != gap and _is_letter_head(out[-1][0])):
            out[-1][0] += s
            changed = True
            if t == gap:
                continue
            if unicodedata.combining(t) and out[-1][1] != gap and _is_letter_head(out[-1][1]):
                out[-1][1] += t
                continue
            s = gap
        elif (t != gap and unicodedata.combining(t) and out and out[-1][1] != gap and _is_letter_head(out[-1][1])):
            out[-1][1] += t
            changed = True
            if s == gap:
                continue
            t = gap
        out.append([s, t])
    return [tuple(p) for p in out] if changed else pairs


def confusion_pairs(pairs, same, gap=GAP_ELEMENT):
    """alignment.py:243-259: every gap-free pair, extended to the LEFT over preceding pairs that hold a gap (so
    that insertions/deletions show as multi-character entries), minus pairs ``same`` calls equal."""
    found = []
    for pos, pair in enumerate(pairs):
        if gap in pair:
            continue
        while pos and gap in pairs[pos - 1]:
            pos -= 1
            pair = tuple((a or '') + (b or '') for a, b in zip(pairs[pos], pair))
        if same is not None and same(*pair):
            continue
        found.append(pair)
    return found


def adjusted_distance(source, target, normalization=None, gt_level=1, want_confusion=False, rules=None):
    """``Alignment.get_adjusted_distance`` (alignment.py:286-356) as a pure function.
    Returns (distance, alignment length, alignment pairs, confusion pairs or None)."""
    rules = rules or _RULES
    classes_on = normalization == 'historic_latin' and gt_level == 1
    src = normalize(source, normalization, gt_level, rules)
    tgt = normalize(target, normalization, gt_level, rules)
    pairs = align_pairs(src, tgt)
    if not isinstance(src, list):
        pairs = recombine_graphemes(pairs)
    same = lambda x, y: _equivalent(x, y, classes_on, rules)     # noqa: E731
    conf = confusion_pairs(pairs, same) if want_confusion else None
    dist = 0.0
    for s, t in pairs:
        if not (s == t or same(s, t)):
            dist += 1.0
    return dist, len(pairs), pairs, conf


class Alignment:
    """Drop-in for ``ocrd_cor_asv_ann.lib.alignment.Alignment`` (the subset ``evaluate`` and the CLIs use)."""

    def __init__(self, gap_element=0, logger=None, confusion=False):
        self.confusion = dict() if confusion else None
        self.gap_element = gap_element
        self.logger = logger or logging.getLogger(__name__)
        self.source_text, self.target_text = [], []

    def set_seqs(self, source_text, target_text):
        self.source_text, self.target_text = source_text, target_text

    def is_bad(self):
        sm = difflib.SequenceMatcher(isjunk=None, autojunk=False)
        sm.set_seqs(self.source_text, self.target_text)
        return bool(sm.quick_ratio() < 0.5 and len(self.source_text) > 5)

    def count_confusion(self, pairs):
        for pair in pairs:
            self.confusion[pair] = self.confusion.get(pair, 0) + 1

    def get_best_alignment(self, eq=None):
        pairs = align_pairs(self.source_text, self.target_text, self.gap_element)
        if not isinstance(self.source_text, list):
            pairs = recombine_graphemes(pairs, self.gap_element)
        if self.confusion is not None:
            self.count_confusion(confusion_pairs(pairs, eq, self.gap_element))
        return pairs

    def get_adjusted_distance(self, source_text, target_text, normalization=None, gtlevel=1, return_alignment=False):
        dist, length, pairs, conf = adjusted_distance(source_text, target_text, normalization, gtlevel,
                                                      want_confusion=self.confusion is not None)
        self.source_text, self.target_text = source_text, target_text
        if conf:
            self.count_confusion(conf)
        return (dist, length, pairs) if return_alignment else (dist, length)

    def get_confusion(self, limit=None):
        """The ``limit`` most frequent non-identity pairs as (count, pair), most frequent first, first seen first
        among equal counts (alignment.py:263-279), and the total number of counted pairs."""
        if self.confusion is None:
            raise Exception("aligner was not configured to count confusion")
        counts, rows, total = [], [], 0         # counts kept negated and ascending for bisect
        for pair, count in self.confusion.items():
            total += count
            if pair[0] == pair[1]:
                continue
            n = len(rows)
            idx = bisect.bisect_left(counts, -count, hi=min(limit or n, n))
            if limit and idx >= limit:
                continue
            counts.insert(idx, -count)
            rows.insert(idx, (count, pair))
        if limit:
            rows = rows[:limit]
        return rows, total

    @staticmethod
    def best_alignment(source_text, target_text, with_confusion=False):
        aligner = Alignment(confusion=with_confusion)
        aligner.set_seqs(source_text, target_text)
        if with_confusion:
            return aligner.get_best_alignment(), aligner.get_confusion()
        return aligner.get_best_alignment()


class Edits:
    """Running error-rate statistics (alignment.py:366-459): length-weighted mean and variance merged with the
    pairwise update of Chan et al. (1979), token histograms of both sides, the worst 1 % of the examples."""

    def __init__(self, logger=None, histogram=False):
        self.logger = logger or logging.getLogger(__name__)
        self.length = 0
        self.mean = self.varia = 0
        self.score = 0
        self.steps = 0
        self.hist1 = {'': 0} if histogram else {}
        self.hist2 = {'': 0} if histogram else {}
        self.worst = []         # (-(mean * length), serial, mean, length, name): ascending = worst first

    def __repr__(self):
        return 'N=%d µ=%.2f σ²=%.2f' % (self.length, self.mean, self.varia)

    def hist(self):
        return {key: (self.hist1.get(key, 0), self.hist2.get(key, 0)) for key in sorted(set(self.hist1) | set(self.hist2))}

    def update(self, steps, length, mean, varia, hist1, hist2):
        if length < 1:
            return
        self.steps += steps
        total = length + self.length
        delta = mean - self.mean
        self.mean = (length * mean + self.length * self.mean) / total
        self.varia = (length * varia + self.length * self.varia + delta ** 2 * length * self.length / total)
        self.length = total
        self.varia /= self.length
        for tok, n in hist1.items():
            self.hist1[tok] = n + self.hist1.get(tok, 0)
        for tok, n in hist2.items():
            self.hist2[tok] = n + self.hist2.get(tok, 0)

    def add(self, dist, length, seq1, seq2, name=None):
        hist1, hist2 = {}, {}
        if self.hist1:
            for tok in seq1:
                hist1[tok] = hist1.get(tok, 0) + 1
        if self.hist2:
            for tok in seq2:
                hist2[tok] = hist2.get(tok, 0) + 1
        rate = dist / length if length else 0
        self.update(1, length, rate, 0, hist1, hist2)
        # insort_left of the reference: a new example goes in front of equally bad ones
        key = -(rate * length)
        pos = bisect.bisect_left([w[0] for w in self.worst], key)
        self.worst.insert(pos, (key, rate, length, name))
        del self.worst[max(int(self.steps * 0.01), 10):]

    def merge(self, edits, name_prefix=None):
        self.update(edits.steps, edits.length, edits.mean, edits.varia, edits.hist1, edits.hist2)
        theirs = [(k, r, n, (name_prefix + (nm or '')) if name_prefix else nm) for k, r, n, nm in edits.worst]
        self.worst = sorted(self.worst + theirs, key=lambda w: w[0])[:max(int(self.steps * 0.01), 10)]


# ----------------------------------------------------------------------------- words (UAX #29)

_MIDNUMLET = set(".\u2018\u2019\u2024\ufe52\uff07\uff0e")
_MIDLETTER = set(":\u00b7\u0387\u055f\u05f4\u2027\ufe13\ufe55\uff1a")
_MIDNUM = set(",;\u037e\u0589\u060c\u060d\u066c\u07f8\u2044\ufe10\ufe14\ufe50\ufe54\uff0c\uff1b")
_NEWLINE = set("\u000b\u000c\u0085\u2028\u2029")
_WSEG = set(" \u1680\u2000\u2001\u2002\u2003\u2004\u2005\u2006\u2008\u2009\u200a\u205f\u3000")
_NOT_ALETTER = ((0x0e00, 0x0eff), (0x1000, 0x109f), (0x1780, 0x17ff), (0x3040, 0x309f), (0x3400, 0x4dbf),
                (0x4e00, 0x9fff), (0xf900, 0xfaff), (0x20000, 0x323af))     # Thai/Lao, Myanmar, Khmer, Hiragana, Han


def _word_break_class(c):
    o = ord(c)
    if c == '\r':
        return 'CR'
    if c == '\n':
        return 'LF'
    if c in _NEWLINE:
        return 'Newline'
    if c == '\u200d':
        return 'ZWJ'
    if c == "'":
        return 'SQ'
    if c == '"':
        return 'DQ'
    if c in _MIDNUMLET:
        return 'MidNumLet'
    if c in _MIDLETTER:
        return 'MidLetter'
    if c in _MIDNUM:
        return 'MidNum'
    if c in _WSEG:
        return 'WSeg'
    if 0x1f1e6 <= o <= 0x1f1ff:
        return 'RI'
    cat = unicodedata.category(c)
    if cat in ('Mn', 'Me', 'Mc') or c == '\u200c':
        return 'Extend'
    if cat == 'Cf':
        return 'Format'
    if cat == 'Nd' or c == '\u066b':
        return 'Numeric'
    if cat == 'Pc' or c == '\u202f':
        return 'ExtendNumLet'
    if 0x30a0 <= o <= 0x30ff or 0x31f0 <= o <= 0x31ff or 0xff66 <= o <= 0xff9f:
        return 'Katakana'
    if 0x05d0 <= o <= 0x05ea or 0x05ef <= o <= 0x05f2 or 0xfb1d <= o <= 0xfb4f:
        return 'Hebrew' if cat == 'Lo' else 'Other'
    if cat[0] == 'L' or cat == 'Nl':
        for lo, hi in _NOT_ALETTER:
            if lo <= o <= hi:
                return 'Other'
        return 'ALetter'
    return 'Other'


def uax29_words(text):
    """Segments of ``text`` between UAX #29 word boundaries (rules WB3-WB999, WB4 = Extend/Format/ZWJ are
    transparent)."""
    n = len(text)
    if not n:
        return
    cls = [_word_break_class(c) for c in text]
    AH = ('ALetter', 'Hebrew')
    skip = ('Extend', 'Format', 'ZWJ')

    def prev_real(i):               # index of the last non-transparent class at or before i (WB4), or -1
        while i >= 0 and cls[i] in skip:
            i -= 1
        return i

    def next_real(i):
        while i < n and cls[i] in skip:
            i += 1
        return i

    start = 0
    for i in range(1, n):
        a, b = cls[i - 1], cls[i]
        brk = True
        if a == 'CR' and b == 'LF':
            brk = False                                         # WB3
        elif a in ('CR', 'LF', 'Newline') or b in ('CR', 'LF', 'Newline'):
            brk = True                                          # WB3a, WB3b
        elif a == 'ZWJ' and unicodedata.category(text[i]) == 'So' and ord(text[i]) >= 0x1f000:
            brk = False                                         # WB3c (Extended_Pictographic, approximated)
        elif a == 'WSeg' and b == 'WSeg':
            brk = False                                         # WB3d
        elif b in skip:
            brk = False                                         # WB4
        else:
            p = prev_real(i - 1)
            A = cls[p] if p >= 0 else 'sot'
            pp = prev_real(p - 1) if p > 0 else -1
            AA = cls[pp] if pp >= 0 else 'sot'
            q = next_real(i + 1)
            C = cls[q] if q < n else 'eot'
            if A in AH and b in AH:
                brk = False                                     # WB5
            elif A in AH and b in ('MidLetter', 'MidNumLet', 'SQ') and C in AH:
                brk = False                                     # WB6
            elif AA in AH and A in ('MidLetter', 'MidNumLet', 'SQ') and b in AH:
                brk = False                                     # WB7
            elif A == 'Hebrew' and b == 'SQ':
                brk = False                                     # WB7a
            elif A == 'Hebrew' and b == 'DQ' and C == 'Hebrew':
                brk = False                                     # WB7b
            elif AA == 'Hebrew' and A == 'DQ' and b == 'Hebrew':
                brk = False                                     # WB7c
            elif A == 'Numeric' and b == 'Numeric':
                brk = False                                     # WB8
            elif A in AH and b == 'Numeric':
                brk = False                                     # WB9
            elif A == 'Numeric' and b in AH:
                brk = False                                     # WB10
            elif AA == 'Numeric' and A in ('MidNum', 'MidNumLet', 'SQ') and b == 'Numeric':
                brk = False                                     # WB11
            elif A == 'Numeric' and b in ('MidNum', 'MidNumLet', 'SQ') and C == 'Numeric':
                brk = False                                     # WB12
            elif A == 'Katakana' and b == 'Katakana':
                brk = False                                     # WB13
            elif A in AH + ('Numeric', 'Katakana', 'ExtendNumLet') and b == 'ExtendNumLet':
                brk = False                                     # WB13a
            elif A == 'ExtendNumLet' and b in AH + ('Numeric', 'Katakana'):
                brk = False                                     # WB13b
            elif A == 'RI' and b == 'RI':
                k, run = p, 0                                   # WB15/16: pair up regional indicators
                while k >= 0 and cls[k] == 'RI':
                    run += 1
                    k = prev_real(k - 1)
                brk = run % 2 == 0
        if brk:
            yield text[start:i]
            start = i
    yield text[start:]


def _unwanted(c):
    sub = unicodedata.category(c)
    return sub[0] in "OMPZS" or sub in ("Cc", "Cf")


def splitwords(text):
    """alignment.py:461-485: UAX #29 words, minus those made only of whitespace, punctuation, marks, symbols or
    control/format characters."""
    try:
        from uniseg.wordbreak import words as segments
    except ImportError:
        segments = uax29_words
    return [w for w in segments(text) if not all(_unwanted(c) for c in w)]


# ----------------------------------------------------------------------------- one line of evaluate()

def line_counts(job):
    """All six alignments of one evaluated line (seq2seq.py:712-733).  ``job`` = (source, target, greedy, beamed,
    normalization, gt_level, want_confusion, rules or None); ``rules`` carries the normalisation state for the
    caller that owns it (the first line of a process), worker processes pass None and use a consumed table.
    Returns {'c_origin': (dist, length, seq1, seq2, confusion pairs), ..., 'w_beamed': ...}."""
    source, target, greedy, beamed, normalization, gt_level, want_confusion, rules = job
    if rules is None:
        rules = _WORKER_RULES
    out = {}
    for name, text in (('origin', source), ('greedy', greedy), ('beamed', beamed)):
        d, n, _, conf = adjusted_distance(text, target, normalization, gt_level, want_confusion, rules)
        out['c_' + name] = (d, n, text, target, conf)
    t_target = splitwords(target)
    tokens = {'greedy': splitwords(greedy), 'beamed': splitwords(beamed), 'origin': splitwords(source)}
    for name in ('origin', 'greedy', 'beamed'):
        d, n, _, _ = adjusted_distance(tokens[name], t_target, normalization, gt_level, False, rules)
        out['w_' + name] = (d, n, tokens[name], t_target, None)
    return out


_WORKER_RULES = HistoricLatinRules()
_WORKER_RULES.single_pending = False


class LineScorer:
    """Maps ``line_counts`` over the lines of a batch on ``workers`` host processes (0 = in this process), keeping
    line order.  The very first line of the process is always scored here, with the stateful rules, because
    the reference's single-codepoint table is consumed by its first string (module docstring)."""

    def __init__(self, workers=0):
        self.pool = None
        if workers and workers > 1:
            import concurrent.futures
            import multiprocessing
            self.pool = concurrent.futures.ProcessPoolExecutor(workers, mp_context=multiprocessing.get_context("fork"))

    def submit(self, jobs):
        """jobs: list of (source, target, greedy, beamed, normalization, gt_level, want_confusion).
        Returns a callable that yields the results in order."""
        jobs = list(jobs)
        head = []
        if jobs and _RULES.single_pending:
            head = [line_counts(jobs[0] + (_RULES,))]
            jobs = jobs[1:]
        if self.pool is None or len(jobs) < 8:
            rest = [line_counts(j + (None,)) for j in jobs]
            return lambda: head + rest
        chunk = max(1, len(jobs) // (4 * self.pool._max_workers))
        it = self.pool.map(line_counts, [j + (None,) for j in jobs], chunksize=chunk)
        return lambda: head + list(it)

    def close(self):
        if self.pool is not None:
            self.pool.shutdown()
            self.pool = None
