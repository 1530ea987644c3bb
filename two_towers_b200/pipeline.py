"""Host input pipeline: text -> token ids -> pinned ring -> device.

The reference's Datasets build a float [T,300] tensor per text on the CPU (one `word2vec[word]` lookup per word) and the
training loop copies those floats to the GPU (simple_two_tower.py:88-110,229-232; enhanced_two_tower.py:144-166;
margin_two_tower.py:96-153).  Here the host produces only the ids - `Vocabulary.encode_batch` runs the native
multi-threaded tokenizer `tt_tokenize_batch` (csrc/pipeline.cu) - and the models gather the embeddings on the device
(`model.attach_embeddings(table, pad_id)`).  For every text, `table[ids]` equals the reference's `text_to_embedding`
output bit for bit when `table` holds the word2vec vectors plus a zero row at `pad_id` (tests/test_pipeline.py).
"""
import ctypes as C
import re

import numpy as np
import torch

from . import _lib as L


def _blob(strings):
    enc = [s.encode("utf-8") for s in strings]
    offs = np.zeros(len(enc) + 1, dtype=np.int64)
    if enc:
        np.cumsum([len(b) for b in enc], out=offs[1:])
    return b"".join(enc), offs


class Vocabulary:
    """word -> row of the embedding table.  `words[i]` maps to id `first_id + i`; row `pad_id` of the table must read
    as zeros (the reference pads with zero vectors).  Default layout: pad row 0, words from row 1."""

    def __init__(self, words, pad_id=0, first_id=1):
        words = list(words)
        self.pad_id, self.first_id = int(pad_id), int(first_id)
        if self.first_id <= self.pad_id < self.first_id + len(words):
            raise ValueError("pad_id collides with a word id")
        self.words = words
        self.index = {}
        for i, w in enumerate(words):
            self.index.setdefault(w, self.first_id + i)
        blob, offs = _blob(words)
        ids = np.arange(self.first_id, self.first_id + len(words), dtype=np.int32)
        self._h = L.lib().tt_vocab_create(blob, offs.ctypes.data, len(words), ids.ctypes.data)
        if not self._h:
            raise RuntimeError("tt_vocab_create failed: " + L.lib().tt_last_error().decode())

    @classmethod
    def from_word2vec(cls, kv, pad_id=0):
        """kv: gensim KeyedVectors (`index_to_key`, `vectors`) as loaded at simple_two_tower.py:167.  Returns
        (vocabulary, table [V+1,E] fp32 CPU tensor with a zero row at pad_id = 0)."""
        if pad_id != 0:
            raise ValueError("from_word2vec puts the pad row at 0")
        vocab = cls(list(kv.index_to_key), pad_id=0, first_id=1)
        vec = torch.as_tensor(np.asarray(kv.vectors), dtype=torch.float32)
        table = torch.cat([torch.zeros(1, vec.shape[1]), vec], 0)
        return vocab, table

    def __len__(self):
        return len(self.index)

    def __contains__(self, w):
        return w in self.index

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                L.lib().tt_vocab_destroy(h)
            except Exception:
                pass

    def table_rows(self):
        """rows the embedding table needs (largest id + 1)."""
        return max(self.pad_id, self.first_id + len(self.words) - 1) + 1

    def encode_batch(self, texts, max_length, out=None, lengths=None, mode=0, threads=0):
        """texts -> int64 ids [N, max_length] (CPU; written into `out` if given, e.g. a pinned ring slot) following
        SimpleDataset.text_to_embedding: lower(), split(), first max_length words, unknown words skipped, pad at the end.
        `lengths` (int32 [N], optional) receives the number of real tokens per row."""
        n = len(texts)
        if out is None:
            out = torch.empty(n, max_length, dtype=torch.int64)
        if out.device.type != "cpu" or not out.is_contiguous() or tuple(out.shape) != (n, max_length) or \
                out.dtype not in (torch.int64, torch.int32):
            raise ValueError("out must be a contiguous CPU int32/int64 tensor of shape [len(texts), max_length]")
        if lengths is not None and (lengths.dtype != torch.int32 or lengths.numel() != n or not lengths.is_contiguous()):
            raise ValueError("lengths must be a contiguous int32 tensor with one entry per text")
        if n == 0:
            return out
        blob, offs = _blob([t.lower() for t in texts])
        L.check(L.lib().tt_tokenize_batch(self._h, blob, offs.ctypes.data, n, max_length, self.pad_id, mode,
                                          1 if out.dtype == torch.int64 else 0, out.data_ptr(),
                                          lengths.data_ptr() if lengths is not None else None, threads), "tt_tokenize_batch")
        return out

    # ---- margin_two_tower.SimpleDataset.text_to_embedding (margin_two_tower.py:96-153): regex rewrite + double lookup.
    # Python-level (the rewrite is Python `re` semantics); ids only, same order of appended vectors as the reference.
    _MARGIN_SUBS = [(re.compile(r'\b(is|are|refers?\s+to)\s+(?:a|an|the)\b'), 'IS'),
                    (re.compile(r'\b(contains?|has|have|includes?)\b'), 'HAS'),
                    (re.compile(r'\b(part|component|element)\s+of\b'), 'PART_OF'),
                    (re.compile(r'\b(controls?|regulates?|manages?)\b'), 'CONTROLS'),
                    (re.compile(r'\b(functions?|works?|operates?)\b'), 'FUNCTIONS'),
                    (re.compile(r'(\d+(?:\.\d+)?)\s*([a-zA-Z]+)'), r'\1_\2')]

    def encode_margin(self, text, max_length):
        original = text.lower().split()
        for rx, rep in self._MARGIN_SUBS:
            text = rx.sub(rep, text.lower())          # each rule sees the lower-cased result of the previous one
        ids = []
        for i, word in enumerate(text.split()):
            if i < len(original) and original[i] in self.index:
                ids.append(self.index[original[i]])
            if word != original[i] and word in self.index:     # IndexError past the original length, as in the reference
                ids.append(self.index[word])
        ids = ids[:max_length]
        return ids + [self.pad_id] * (max_length - len(ids))

    def encode_margin_batch(self, texts, max_length):
        return torch.tensor([self.encode_margin(t, max_length) for t in texts], dtype=torch.int64).reshape(len(texts), max_length)


class PinnedIdRing:
    """`depth` pinned host slots + device slots for (query ids, doc ids) batches; `put` tokenises straight into a pinned
    slot and starts the host->device copy on a private copy stream, `get` makes the compute stream wait for it.
    The training loop stays one batch ahead: the copy of batch i+1 overlaps the step on batch i."""

    def __init__(self, vocab, batch_size, query_len, doc_len, device, depth=2, dtype=torch.int64):
        self.vocab, self.device, self.depth = vocab, torch.device(device), depth
        if self.device.type != "cuda":
            raise RuntimeError("PinnedIdRing feeds CUDA models (no CPU fallback)")
        mk = lambda T: [torch.empty(batch_size, T, dtype=dtype).pin_memory() for _ in range(depth)]
        self.hq, self.hd = mk(query_len), mk(doc_len)
        self.dq = [torch.empty(batch_size, query_len, dtype=dtype, device=self.device) for _ in range(depth)]
        self.dd = [torch.empty(batch_size, doc_len, dtype=dtype, device=self.device) for _ in range(depth)]
        self.copy_stream = torch.cuda.Stream(self.device)
        self.ready = [torch.cuda.Event() for _ in range(depth)]
        self.free = [None] * depth                 # event: the step that last read slot i has been queued and finished
        self.head = self.tail = 0

    def put(self, queries, docs):
        i = self.head % self.depth
        n = len(queries)
        if self.free[i] is not None:
            self.free[i].synchronize()             # host must not overwrite a pinned slot still being copied / read
        self.vocab.encode_batch(queries, self.hq[i].shape[1], out=self.hq[i][:n])
        self.vocab.encode_batch(docs, self.hd[i].shape[1], out=self.hd[i][:n])
        with torch.cuda.stream(self.copy_stream):
            if self.free[i] is not None:
                self.copy_stream.wait_event(self.free[i])
            self.dq[i][:n].copy_(self.hq[i][:n], non_blocking=True)
            self.dd[i][:n].copy_(self.hd[i][:n], non_blocking=True)
            self.ready[i].record(self.copy_stream)
        self.head += 1
        return n

    def get(self, n=None):
        """Device id tensors of the oldest pending batch; the current stream waits for its copy."""
        if self.tail >= self.head:
            raise RuntimeError("PinnedIdRing.get() without a pending put()")
        i = self.tail % self.depth
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self.ready[i])
        self.tail += 1
        q, d = (self.dq[i], self.dd[i]) if n is None else (self.dq[i][:n], self.dd[i][:n])
        return q, d, i

    def release(self, slot):
        """Call after the step that consumed `slot` has been queued on the current stream."""
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.free[slot] = ev


def id_batches(vocab, queries, docs, batch_size, query_len, doc_len, device, drop_last=False):
    """Generator of (query_ids, doc_ids) device tensors over aligned text lists, one batch ahead of the consumer."""
    ring = PinnedIdRing(vocab, batch_size, query_len, doc_len, device)
    starts = list(range(0, len(queries), batch_size))
    if drop_last and starts and len(queries) - starts[-1] < batch_size:
        starts.pop()
    sizes = []
    for k, s in enumerate(starts[:1]):
        sizes.append(ring.put(queries[s:s + batch_size], docs[s:s + batch_size]))
    for k, s in enumerate(starts):
        if k + 1 < len(starts):
            s2 = starts[k + 1]
            sizes.append(ring.put(queries[s2:s2 + batch_size], docs[s2:s2 + batch_size]))
        q, d, slot = ring.get(sizes[k])
        yield q, d
        ring.release(slot)
