"""BLEU over decoded captions (the caller side of decode).

The reference scores each decoded caption against the image's COCO captions with NLTK:
`sentence_bleu(references, caption_split, weights=w, smoothing_function=chencherry.method0)` for eight weight
vectors (AttemptFour/evaluate.py:178-226), and loops `sentence_bleu` in a Cython helper
(CNN_RNN/cython/cython_functions.pyx:24-50).  NLTK is not installed in this image, so its algorithm
(nltk/translate/bleu_score.py, NLTK 3.x: modified n-gram precision with clipping, closest reference length, brevity
penalty, geometric mean, SmoothingFunction.method0) is restated here in plain Python; host-side text work, no GPU.
"""
import math
import sys
from collections import Counter
from fractions import Fraction

# evaluate.py:183-192: individual 1..4-gram scores, then cumulative BLEU-1..4
EVALUATE_WEIGHTS = [(1, 0, 0, 0), (0, 1, 0, 0), (0, 0, 1, 0), (0, 0, 0, 1),
                    (1.0, 0, 0, 0), (1.0 / 2, 1.0 / 2, 0, 0), (1.0 / 3, 1.0 / 3, 1.0 / 3, 0),
                    (1.0 / 4, 1.0 / 4, 1.0 / 4, 1.0 / 4)]


def _ngrams(seq, n):
    return [tuple(seq[i:i + n]) for i in range(len(seq) - n + 1)]


def modified_precision(references, hypothesis, n):
    """(clipped matches, total hypothesis n-grams) -- bleu_score.modified_precision."""
    counts = Counter(_ngrams(hypothesis, n)) if len(hypothesis) >= n else Counter()
    max_counts = {}
    for ref in references:
        rc = Counter(_ngrams(ref, n)) if len(ref) >= n else Counter()
        for g in counts:
            max_counts[g] = max(max_counts.get(g, 0), rc[g])
    clipped = {g: min(c, max_counts[g]) for g, c in counts.items()}
    return sum(clipped.values()), max(1, sum(counts.values()))


def closest_ref_length(references, hyp_len):
    return min((len(r) for r in references), key=lambda rl: (abs(rl - hyp_len), rl))


def brevity_penalty(closest_ref_len, hyp_len):
    if hyp_len > closest_ref_len:
        return 1.0
    if hyp_len == 0:
        return 0.0
    return math.exp(1 - closest_ref_len / hyp_len)


def corpus_bleu(list_of_references, hypotheses, weights=(0.25, 0.25, 0.25, 0.25)):
    """bleu_score.corpus_bleu with SmoothingFunction().method0 (no smoothing: a zero precision becomes
    sys.float_info.min, i.e. the score collapses to ~0)."""
    assert len(list_of_references) == len(hypotheses)
    num, den = Counter(), Counter()
    hyp_lengths = ref_lengths = 0
    for references, hypothesis in zip(list_of_references, hypotheses):
        for i in range(1, len(weights) + 1):
            a, b = modified_precision(references, hypothesis, i)
            num[i] += a
            den[i] += b
        hyp_lengths += len(hypothesis)
        ref_lengths += closest_ref_length(references, len(hypothesis))
    bp = brevity_penalty(ref_lengths, hyp_lengths)
    if num[1] == 0:
        return 0.0
    p_n = [Fraction(num[i], den[i]) if num[i] != 0 else sys.float_info.min for i in range(1, len(weights) + 1)]
    s = math.fsum(w * math.log(p) for w, p in zip(weights, p_n) if p > 0)
    return bp * math.exp(s)


def sentence_bleu(references, hypothesis, weights=(0.25, 0.25, 0.25, 0.25)):
    return corpus_bleu([references], [hypothesis], weights)


def ids_to_words(ids, index_word, end_token="<end>", pad_id=0):
    """Token ids of one caption -> words, cut at the first <end> / pad (tokenizer.sequences_to_texts + the '<pad>'
    removal of evaluate.py:208-212)."""
    out = []
    for i in ids:
        i = int(i)
        if i == pad_id:
            break
        w = index_word.get(i, index_word.get(str(i), "<unk>"))
        if w == end_token:
            break
        out.append(w)
    return out


def compute_bleu(caption_words, references):
    """evaluate.compute_bleu: the eight scores of one candidate against its reference captions (lists of words)."""
    return [sentence_bleu(references, caption_words, weights=w) for w in EVALUATE_WEIGHTS]
