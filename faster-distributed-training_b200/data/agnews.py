"""AG_NEWS text-classification pipeline.

The reference streams torchtext's ``AG_NEWS`` datapipe, cleans each sentence (strip HTML, remove
URLs, gensim stop-word removal), tokenises the batch with HF ``bert-base-uncased``
(``padding='longest'``) inside the collate function and yields
``(input_ids, label 1..4, token_type_ids, attention_mask)`` (``transformer_test.py:62-149``).

torchtext and gensim are not installable here and the GPU box has no network, so:

* the corpus is read from the standard ``train.csv`` / ``test.csv`` files
  (``"class","title","description"``) if they exist under ``root``; otherwise
  :class:`fdt.data.synthetic.SyntheticAGNews` stands in;
* cleaning uses the same two regular expressions and a built-in English stop-word list;
* the tokenizer is ``bert-base-uncased`` through ``transformers`` when its files are cached
  locally, else :class:`HashWordTokenizer` — a deterministic lower-cased word/punctuation
  tokenizer hashed into the same 30522-entry id space with the BERT special ids
  ([PAD]=0, [UNK]=100, [CLS]=101, [SEP]=102), so shapes, dtypes and vocabulary size match.
"""
from __future__ import annotations

import csv
import os
import re
import zlib
from typing import List, Optional, Sequence, Tuple

import torch

from .synthetic import SyntheticAGNews

VOCAB_SIZE = 30522  # bert-base-uncased, transformer_test.py:141-143

_HTML = re.compile(r"<.*?>")
_URL = re.compile(r"\s*(?:https?://)?www\.\S*\.[A-Za-z]{2,5}\s*")
_WORD = re.compile(r"[a-z0-9]+|[^\sa-z0-9]")

STOPWORDS = frozenset("""a about above after again against all am an and any are aren't as at be because been
before being below between both but by can't cannot could couldn't did didn't do does doesn't doing don't down
during each few for from further had hadn't has hasn't have haven't having he he'd he'll he's her here here's hers
herself him himself his how how's i i'd i'll i'm i've if in into is isn't it it's its itself let's me more most
mustn't my myself no nor not of off on once only or other ought our ours ourselves out over own same shan't she
she'd she'll she's should shouldn't so some such than that that's the their theirs them themselves then there
there's these they they'd they'll they're they've this those through to too under until up very was wasn't we we'd
we'll we're we've were weren't what what's when when's where where's which while who who's whom why why's with
won't would wouldn't you you'd you'll you're you've your yours yourself yourselves""".split())


def striphtml(data: str) -> str:
    return _HTML.sub("", data)


def remove_url(data: str) -> str:
    return _URL.sub(" ", data).strip()


def remove_stopwords(data: str) -> str:
    return " ".join(w for w in data.split() if w.lower() not in STOPWORDS)


def clean(data: str) -> str:
    return remove_stopwords(remove_url(striphtml(data)))


class HashWordTokenizer:
    PAD, UNK, CLS, SEP = 0, 100, 101, 102
    vocab_size = VOCAB_SIZE

    def __init__(self, max_len: int = 512):
        self.max_len = max_len

    def encode(self, text: str) -> List[int]:
        ids = [self.CLS]
        for w in _WORD.findall(text.lower())[: self.max_len - 2]:
            ids.append(1000 + zlib.crc32(w.encode()) % (self.vocab_size - 1000))
        ids.append(self.SEP)
        return ids

    def __call__(self, sentences: Sequence[str], padding="longest", return_tensors="pt", pad_to: Optional[int] = None):
        enc = [self.encode(s) for s in sentences]
        L = pad_to or max(len(e) for e in enc)
        ids = torch.zeros(len(enc), L, dtype=torch.long)
        mask = torch.zeros(len(enc), L, dtype=torch.long)
        for i, e in enumerate(enc):
            e = e[:L]
            ids[i, : len(e)] = torch.tensor(e)
            mask[i, : len(e)] = 1
        return {"input_ids": ids, "token_type_ids": torch.zeros_like(ids), "attention_mask": mask}


def get_tokenizer():
    try:
        from transformers import AutoTokenizer
        return AutoTokenizer.from_pretrained("bert-base-uncased", local_files_only=True)
    except Exception:
        return HashWordTokenizer()


def get_tokenizer_size() -> int:
    return getattr(get_tokenizer(), "vocab_size", VOCAB_SIZE)


class AGNewsCSV(torch.utils.data.Dataset):
    def __init__(self, path: str):
        self.rows: List[Tuple[int, str]] = []
        with open(path, newline="", encoding="utf-8") as f:
            for r in csv.reader(f):
                if len(r) >= 3 and r[0].strip().isdigit():
                    self.rows.append((int(r[0]), r[1] + " " + r[2]))

    def __len__(self):
        return len(self.rows)

    def __getitem__(self, i):
        return self.rows[i]


class AGNewsDataset:
    """``AG_NEWS_DATASET`` equivalent: holds the splits, the tokenizer and the collate function."""

    def __init__(self, tokenizer=None, batch_size: int = 32, root: str = "./data", synthetic: bool = False,
                 seq_len: int = 128, n_train: int = 120000, n_test: int = 7600):
        self.tokenizer = tokenizer or get_tokenizer()
        self.batch_size = batch_size
        tr, te = os.path.join(root, "ag_news", "train.csv"), os.path.join(root, "ag_news", "test.csv")
        self.synthetic = synthetic or not (os.path.isfile(tr) and os.path.isfile(te))
        if self.synthetic:
            self.train_ds = SyntheticAGNews(n_train, seq_len, seed=0)
            self.test_ds = SyntheticAGNews(n_test, seq_len, seed=1)
        else:
            self.train_ds, self.test_ds = AGNewsCSV(tr), AGNewsCSV(te)

    def generate_batch(self, data_batch):
        labels, raw = zip(*data_batch)
        tmp = self.tokenizer([clean(x) for x in raw], padding="longest", return_tensors="pt")
        return (tmp["input_ids"], torch.tensor(labels, dtype=torch.long), tmp["token_type_ids"], tmp["attention_mask"])

    def load_data(self, distributed: bool = False, num_workers: int = 2, rank: int = 0, world: int = 1, device="cpu"):
        """Returns ``(train_loader, test_loader, train_sampler)``."""
        from .loader import BackgroundLoader, TensorBatchLoader
        from ..parallel.sampler import ShardedSampler
        pin = torch.device(device).type == "cuda"
        if self.synthetic:
            samp = ShardedSampler(len(self.train_ds), world if distributed else 1, rank if distributed else 0, True)
            tr = TensorBatchLoader(self.train_ds.tensors(), self.batch_size, samp, drop_last=True, pin=pin)
            tsamp = ShardedSampler(len(self.test_ds), world if distributed else 1, rank if distributed else 0, False)
            te = TensorBatchLoader(self.test_ds.tensors(), self.batch_size, tsamp, shuffle=False, drop_last=False, pin=pin)
            return tr, te, samp
        from torch.utils.data import DataLoader
        samp = ShardedSampler(len(self.train_ds), world if distributed else 1, rank if distributed else 0, True)
        kw = dict(batch_size=self.batch_size, collate_fn=self.generate_batch, num_workers=num_workers,
                  pin_memory=pin, persistent_workers=num_workers > 0)
        tr = BackgroundLoader(DataLoader(self.train_ds, sampler=samp, drop_last=True, **kw))
        te = BackgroundLoader(DataLoader(self.test_ds, shuffle=False, **kw))
        return tr, te, samp
