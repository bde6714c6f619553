"""Tokenise a text corpus into the flat int32 ``.npy`` token file the native loader reads.

    python tools/prepare_tokens.py --tokenizer Qwen/Qwen3-0.6B --out data/wikitext2_train.npy --dataset wikitext --subset wikitext-2-raw-v1
    python tools/prepare_tokens.py --tokenizer ./my_tokenizer --out data/corpus.npy --text_files a.txt b.txt
    torchrun … tools/train.py --native_loader true --token_file data/corpus.npy …

(The reference tokenises inside the training job with ``datasets.map`` on every start, ``scaletorch/data/dataset.py:463-478``.)
"""
from __future__ import annotations

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaletorch_b200.data.dataset import DatasetProcessor, export_flat_tokens  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokenizer", required=True)
    ap.add_argument("--out", required=True)
    ap.add_argument("--dataset", default=None, help="HF dataset name or a load_from_disk directory")
    ap.add_argument("--subset", default=None)
    ap.add_argument("--split", default="train")
    ap.add_argument("--text_column", default="text")
    ap.add_argument("--text_files", nargs="*", default=None, help="plain-text files, one document per line")
    ap.add_argument("--no_eos", action="store_true")
    a = ap.parse_args()
    tok = DatasetProcessor.create_tokenizer(a.tokenizer)
    if a.text_files:
        def lines():
            for f in a.text_files:
                with open(f, encoding="utf-8") as fh:
                    for ln in fh:
                        if ln.strip():
                            yield ln.rstrip("\n")
        texts = lines()
    else:
        ds = DatasetProcessor.load_dataset(a.dataset, a.split, a.subset)
        texts = (row[a.text_column] for row in ds)
    n = export_flat_tokens(texts, tok, a.out, eos_token_id=None if a.no_eos else tok.eos_token_id)
    print(f"wrote {n:,} tokens to {a.out}")


if __name__ == "__main__":
    main()
