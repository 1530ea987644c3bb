"""TEST INFRASTRUCTURE ONLY - generate tests/golden/pipeline.json by running the UNMODIFIED reference Datasets'
text_to_embedding (simple_two_tower.py:88-110, enhanced_two_tower.py:144-166, margin_two_tower.py:96-153) over a fake
word2vec whose vector for word i is (i+1, ...): the first component of every returned row is then the table row id
(0 = zero padding).  Run in the build container:  python oracle/gen_golden_pipeline.py"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.ref_import import load_reference  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "pipeline.json")

WORDS = ["the", "a", "an", "is", "are", "of", "what", "how", "to", "cook", "pasta", "ai", "machine", "learning",
         "capital", "france", "paris", "heart", "has", "have", "contains", "part", "controls", "functions", "works",
         "5_kg", "2.5_mg", "is", "caf\u00e9", "na\u00efve", "\u00fcber", "stra\u00dfe", "10", "kg", "water", "salt",
         "part_of", "IS", "HAS", "refers", "component", "blood", "pumps", "body", "photosynthesis", "light", "energy",
         "nbsp", "em", "line", "ideographic", "ogham", "math", "narrow", "nel", "ok", "i\u0307stanbul", "\u00e7a"]
TEXTS = [
    "what is ai", "How To Cook PASTA", "", "   ", "zzz qqq", "the heart is a part of the body",
    "The heart pumps blood; the heart controls 5 kg of water and 2.5mg salt",
    "tabs\tand\nnewlines\r\nare  whitespace\x0bto\x0csplit\x1con\x1d\x1e\x1f ok",
    "nbsp\u00a0the\u2003em\u2028line\u3000ideographic\u1680ogham\u205fmath\u202fnarrow\u0085nel\u2009a\u200bthe",
    "Caf\u00e9 NA\u00cfVE \u00dcBER Stra\u00dfe the", "\u0130stanbul the \u00c7a",
    "machine learning is a branch of ai " * 12, "unknown " * 40 + "the pasta",
    "photosynthesis refers to the light energy", "it contains water and has salt and includes pasta",
    "a component of the machine functions and works and operates", "10 kg 10kg 3.5 kg of pasta is the capital",
    "the the the the the the the the the the the the", "What  IS   the CAPITAL of FRANCE",
]
MAX_LENGTHS = [1, 5, 10, 30]


class FakeW2V:
    vector_size = 3

    def __init__(self, words):
        self.index = {}
        for i, w in enumerate(words):
            self.index.setdefault(w, i + 1)

    def __getitem__(self, w):
        i = self.index[w]                  # KeyError for unknown words, like gensim
        return np.array([i, 2 * i, -i], dtype=np.float32)

    def __contains__(self, w):
        return w in self.index


def main():
    simple, margin, enhanced = load_reference()
    kv = FakeW2V(WORDS)
    rec = {"words": WORDS, "texts": TEXTS, "max_lengths": MAX_LENGTHS, "simple": {}, "enhanced": {}, "margin": {}}
    for L in MAX_LENGTHS:
        ds_s = simple.SimpleDataset([], [], kv, max_length=L)
        ds_e = enhanced.EnhancedDataset([], [], kv, max_length=L)
        rec["simple"][str(L)] = [ds_s.text_to_embedding(t)[:, 0].long().tolist() for t in TEXTS]
        rec["enhanced"][str(L)] = [ds_e.text_to_embedding(t)[:, 0].long().tolist() for t in TEXTS]
        rec["margin"][str(L)] = [margin.SimpleDataset.text_to_embedding(t, kv, L)[:, 0].long().tolist() for t in TEXTS]
        assert all(len(r) == L for r in rec["simple"][str(L)] + rec["margin"][str(L)])
    json.dump(rec, open(OUT, "w"), ensure_ascii=True, indent=0)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
