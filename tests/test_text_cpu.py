"""Host-side IMDB preparation (pytorch_on_language_distr.py:34-103): rm_tags, the keras pad_sequences call the script
makes, attention masks, and a tokenizer round trip with a corpus-built vocabulary."""
import numpy as np


def test_rm_tags_is_the_references_regex():
    from b200net.text import rm_tags
    assert rm_tags("good<br /><br />movie <b>really</b>") == "goodmovie really"
    assert rm_tags("no tags") == "no tags"


def test_pad_sequences_keras_semantics():
    from b200net.text import pad_sequences, attention_masks
    seqs = [[101, 7, 8, 102], [101, 5, 6, 7, 8, 9, 102], []]
    out = pad_sequences(seqs, maxlen=5, dtype="long", value=0, truncating="post", padding="post")
    assert out.dtype == np.int64 and out.shape == (3, 5)
    assert out.tolist() == [[101, 7, 8, 102, 0], [101, 5, 6, 7, 8], [0, 0, 0, 0, 0]]
    pre = pad_sequences(seqs, maxlen=5)                         # keras defaults: pad and truncate at the front
    assert pre.tolist() == [[0, 101, 7, 8, 102], [6, 7, 8, 9, 102], [0, 0, 0, 0, 0]]
    assert attention_masks(out)[0] == [1, 1, 1, 1, 0]
    assert pad_sequences([[1, 2], [3]]).tolist() == [[1, 2], [0, 3]]


def test_tokenizer_with_corpus_vocabulary(tmp_path):
    from transformers import BertTokenizer
    from b200net.text import build_wordpiece_vocab
    texts = ["The film was great, truly great!", "A dull and boring plot."]
    vocab = build_wordpiece_vocab(texts, str(tmp_path / "vocab.txt"))
    tok = BertTokenizer(vocab=vocab, do_lower_case=True)      # transformers 5: `vocab=`; 4.x: `vocab_file=`
    ids = tok.encode(texts[0], add_special_tokens=True, max_length=16, truncation=True)
    assert ids[0] == tok.cls_token_id == 2 and ids[-1] == tok.sep_token_id == 3 and tok.pad_token_id == 0
    assert tok.unk_token_id not in ids
    assert tok.convert_ids_to_tokens(ids[1:-1]) == ["the", "film", "was", "great", ",", "truly", "great", "!"]
