"""Host-side text preparation of the IMDB script (pytorch_on_language_distr.py:34-103), without the packages the
reference imports for it that are not in this image (keras' `pad_sequences`) -- same names, same semantics -- so the
script's statements run unchanged.  String work stays on the host; the model sees int64 id / mask tensors."""
import re

import numpy as np


def rm_tags(text):
    """pytorch_on_language_distr.py:34-36: strip HTML tags (`<br />` in the IMDB reviews)."""
    re_tag = re.compile(r'<[^>]+>')
    return re_tag.sub('', text)


def pad_sequences(sequences, maxlen=None, dtype="int32", padding="pre", truncating="pre", value=0.0):
    """keras.preprocessing.sequence.pad_sequences as the reference calls it (:74-76:
    `pad_sequences(input_ids, maxlen=MAX_LEN, dtype="long", truncating="post", padding="post")`): every sequence cut
    to `maxlen` (from the front for 'pre', from the back for 'post') and padded with `value` on the chosen side."""
    lengths = [len(s) for s in sequences]
    if maxlen is None:
        maxlen = max(lengths) if lengths else 0
    dt = np.int64 if dtype in ("long", "int64") else np.dtype(dtype)
    out = np.full((len(sequences), maxlen), value, dtype=dt)
    for i, s in enumerate(sequences):
        if not len(s):
            continue
        if truncating == "pre":
            trunc = s[-maxlen:]
        elif truncating == "post":
            trunc = s[:maxlen]
        else:
            raise ValueError('Truncating type "%s" not understood' % truncating)
        trunc = np.asarray(trunc, dtype=dt)
        if padding == "post":
            out[i, :len(trunc)] = trunc
        elif padding == "pre":
            out[i, -len(trunc):] = trunc
        else:
            raise ValueError('Padding type "%s" not understood' % padding)
    return out


def attention_masks(input_ids):
    """:86-96: 1 for real tokens (id > 0), 0 for padding."""
    return [[int(token_id > 0) for token_id in sent] for sent in input_ids]


def build_wordpiece_vocab(texts, path, max_words=20000):
    """A vocab.txt for transformers.BertTokenizer built from the corpus itself (whole lower-cased words), for boxes
    where `BertTokenizer.from_pretrained('bert-base-uncased')` (:60) cannot download its vocabulary.  The five
    special tokens come first with bert-base-uncased's own ids for [PAD] = 0; [UNK], [CLS], [SEP], [MASK] follow."""
    from collections import Counter
    counts = Counter()
    for t in texts:
        counts.update(re.findall(r"[a-z0-9]+|[^\sa-z0-9]", t.lower()))
    words = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + [w for w, _ in counts.most_common(max_words)]
    with open(path, "w") as fh:
        fh.write("\n".join(words) + "\n")
    return path
