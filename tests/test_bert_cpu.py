"""BERT / IMDB path without a GPU: the oracle (transformers on CPU executing the reference's statements,
pytorch_on_language_distr.py:257-275) reproduces the committed golden vectors; the b200net module tree is a drop-in
for transformers' (same state_dict keys, shapes, parameter order, output indexing, optimizer / schedule objects)."""
import os
import warnings

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def bert_golden():
    return torch.load(os.path.join(ROOT, "tests", "golden", "bert_small.pt"), weights_only=False)


def test_oracle_reproduces_golden(bert_golden):
    from oracle import bert_oracle as B
    from oracle.torch_oracle import digest
    g, g64 = bert_golden["case"], bert_golden["case_f64"]
    model = B.build_reference_model(num_layers=bert_golden["layers"])
    res = B.run_case(model, bert_golden["n"], bert_golden["s"], bert_golden["steps"])
    for i in range(bert_golden["steps"]):
        assert abs(res["losses"][i] - g["losses"][i]) <= 1e-5 * abs(g["losses"][i])
        tol = 1e-4 + 3 * ((g["logits"][i] - g64["logits"][i]).abs().max() / g64["logits"][i].abs().max()).item()
        assert ((res["logits"][i] - g["logits"][i]).abs().max() / g["logits"][i].abs().max()).item() <= tol
    assert torch.equal(res["eval_logits"].argmax(1), g["eval_logits"].argmax(1))
    sd = model.state_dict()
    for k, d in g["weights"].items():
        assert torch.allclose(digest(sd[k])["samples"], d["samples"], rtol=1e-4, atol=1e-6), k
    # the fp64 arbiter agrees with fp32 to ~1e-6 on this path (LayerNorm, unlike small-batch BatchNorm, is
    # well conditioned), so the north-star tolerances apply without arbitration
    assert abs(g["losses"][0] - g64["losses"][0]) < 1e-6


def test_synthetic_batch_shape():
    """[CLS] ... [SEP] pad, mask = (id > 0) (pytorch_on_language_distr.py:93)"""
    from oracle import bert_oracle as B
    ids, mask, labels = B.make_batch(8, 256, 0)
    assert ids.shape == (8, 256) and ids.dtype == torch.int64 and labels.shape == (8,)
    assert (ids[:, 0] == 101).all() and torch.equal(mask, (ids > 0).long())
    lens = mask.sum(1)
    assert all(int(ids[i, lens[i] - 1]) == 102 for i in range(8))
    assert int(ids.max()) < 30522 and set(labels.tolist()) <= {0, 1}


def test_module_tree_is_transformers_compatible():
    import transformers
    from b200net import bert
    ours = bert.BertForSequenceClassification(bert.BertConfig(num_labels=2, num_hidden_layers=2))
    ref = transformers.BertForSequenceClassification(transformers.BertConfig(num_labels=2, num_hidden_layers=2))
    so, sr = ours.state_dict(), ref.state_dict()
    assert list(so.keys()) == list(sr.keys())
    assert all(so[k].shape == sr[k].shape and so[k].dtype == sr[k].dtype for k in so)
    assert [n for n, _ in ours.named_parameters()] == [n for n, _ in ref.named_parameters()]
    ours.load_state_dict(sr)
    # a transformers config object is accepted as is
    again = bert.BertForSequenceClassification(ref.config)
    assert again.config.num_hidden_layers == 2 and again.num_labels == 2


def test_bert_base_parameter_count():
    """109 483 778 parameters in 201 tensors for num_labels=2"""
    from b200net import bert
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m = bert.BertForSequenceClassification.from_pretrained("bert-base-uncased", num_labels=2,
                                                               output_attentions=False, output_hidden_states=False)
    ps = list(m.parameters())
    assert len(ps) == 201 and sum(p.numel() for p in ps) == 109483778
    assert len(bert._encoder_params(m)) == 197
    assert float(m.bert.embeddings.word_embeddings.weight[0].abs().sum()) == 0.0    # padding_idx row


def test_output_indexing_protocol():
    """outputs[0] is the loss when labels are given and the logits otherwise (:265, :318, :364)"""
    from b200net.bert import SequenceClassifierOutput
    loss, logits = torch.tensor(0.5), torch.zeros(4, 2)
    o = SequenceClassifierOutput(loss, logits)
    assert o[0] is loss and o[1] is logits and len(o) == 2 and o.loss is loss and o["logits"] is logits
    o = SequenceClassifierOutput(None, logits)
    assert o[0] is logits and len(o) == 1 and tuple(o) == (logits,)


def test_adamw_and_schedule_objects():
    """AdamW(model.parameters(), lr=2e-5, eps=1e-8) and get_linear_schedule_with_warmup(optimizer,
    num_warmup_steps=0, num_training_steps=total) (:167-183)"""
    from b200net import bert
    w = torch.nn.Parameter(torch.zeros(4))
    opt = bert.AdamW([w], lr=2e-5, eps=1e-8)
    assert opt.defaults["weight_decay"] == 0.0 and opt.defaults["eps"] == 1e-8
    assert bert.AdamW([w]).defaults["eps"] == 1e-6
    sched = bert.get_linear_schedule_with_warmup(opt, num_warmup_steps=0, num_training_steps=10)
    opt.step()
    sched.step()
    assert abs(opt.param_groups[0]["lr"] - 2e-5 * 0.9) < 1e-15
    warm = bert.get_linear_schedule_with_warmup(bert.AdamW([w], lr=1.0), 4, 10)
    assert warm.get_last_lr()[0] == 0.0


def test_cpu_input_is_rejected():
    from b200net import bert
    m = bert.BertForSequenceClassification(bert.BertConfig(num_hidden_layers=1))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.zeros(2, 8, dtype=torch.long))
