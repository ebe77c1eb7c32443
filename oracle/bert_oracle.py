"""ORACLE -- TEST INFRASTRUCTURE ONLY (BERT / IMDB path).  Never imported by the product package (b200net); only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.

The reference's language script (pytorch_on_language_distr.py) has no arithmetic of its own: it calls
`transformers.BertForSequenceClassification`, `transformers.AdamW`, `get_linear_schedule_with_warmup` and
`torch.nn.utils.clip_grad_norm_`.  transformers is a third-party dependency that is not vendored under
/root/reference and that the reference does not pin (no requirements file; the script's API -- `AdamW` exported from
transformers, `outputs[0]` tuples -- dates it to transformers 2.x/3.x).  This oracle executes the SAME third-party
model class from the transformers build in this image (5.5) on CPU, statement for statement as the reference does
(:258-275), with the eager attention implementation (the published BertSelfAttention algorithm: softmax(QK^T/8 +
(1-mask)*min) V).  `transformers.AdamW` no longer exists in 5.x; its published update rule (Adam with
correct_bias=True and weight_decay=0.0, the defaults the reference relies on at :167-170) is identical to
torch.optim.AdamW(weight_decay=0), which is what the oracle steps with.

Pinning: the reference stores no replayable golden vector for this path (its printed losses need the pretrained
checkpoint and the IMDB csv).  Parity is therefore pinned to the third-party module: golden logits / loss /
after-step weights produced HERE by oracle/make_golden.py from transformers on CPU (fp32 and fp64) are committed as
tests/golden/bert_small.pt and re-checked on CPU (transformers is present on the GPU box too) and on the GPU
against b200net.
"""
import torch

SEED_WEIGHTS = 42   # seed_val = 42, pytorch_on_language_distr.py:212
SEED_DATA = 0


def make_batch(n, s, seed=SEED_DATA, device="cpu", vocab=30522):
    """Synthetic IMDB-shaped batch: [CLS]=101, word-piece ids, [SEP]=102, pad 0;
    attention mask = (id > 0) as the reference builds it (:93); binary labels."""
    g = torch.Generator().manual_seed(seed)
    ids = torch.zeros(n, s, dtype=torch.int64)
    lo = min(32, max(2, s // 4))
    for i in range(n):
        length = int(torch.randint(lo, s, (1,), generator=g).item())
        ids[i, 0] = 101
        ids[i, 1:length] = torch.randint(1000, vocab, (length - 1,), generator=g)
        ids[i, length] = 102
    mask = (ids > 0).long()
    labels = torch.randint(0, 2, (n,), generator=g)
    return ids.to(device), mask.to(device), labels.to(device)


def build_reference_model(seed=SEED_WEIGHTS, num_layers=12, dropout=0.0, dtype=torch.float32):
    """BertForSequenceClassification(bert-base config, num_labels=2) (:155-161) with random initialisation (no
    checkpoint offline).  dropout=0 makes runs comparable across RNG implementations."""
    import transformers
    torch.manual_seed(seed)
    cfg = transformers.BertConfig(num_labels=2, num_hidden_layers=num_layers, hidden_dropout_prob=dropout,
                                  attention_probs_dropout_prob=dropout)
    try:
        cfg._attn_implementation = "eager"
    except Exception:
        pass
    model = transformers.BertForSequenceClassification(cfg)
    return model.to(dtype)


def make_optimizer(model, total_steps, lr=2e-5, eps=1e-8):
    """AdamW(lr=2e-5, eps=1e-8) (:167-170) + get_linear_schedule_with_warmup(0 warm-up, total_steps) (:181-183)."""
    opt = torch.optim.AdamW(model.parameters(), lr=lr, eps=eps, weight_decay=0.0)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda t: max(0.0, float(total_steps - t) / float(max(1, total_steps))))
    return opt, sched


def train_step(model, optimizer, scheduler, ids, mask, labels, clip=torch.nn.utils.clip_grad_norm_):
    """Body of the reference's training loop, pytorch_on_language_distr.py:257-275."""
    model.zero_grad()
    outputs = model(ids, token_type_ids=None, attention_mask=mask, labels=labels)
    loss = outputs[0]
    loss.backward()
    clip(model.parameters(), 1.0)
    optimizer.step()
    scheduler.step()
    return loss.item(), outputs[1].detach()


def eval_step(model, ids, mask):
    """Validation body, :311-318: logits = outputs[0] when no labels are passed."""
    with torch.no_grad():
        outputs = model(ids, token_type_ids=None, attention_mask=mask)
    return outputs[0]


def run_case(model, n, s, steps, total_steps=30, make_opt=make_optimizer, clip=torch.nn.utils.clip_grad_norm_,
             device="cpu", keep_grads=False):
    """`steps` training steps on one fixed batch, then an eval pass on a second batch."""
    ids, mask, labels = make_batch(n, s, SEED_DATA, device)
    opt, sched = make_opt(model, total_steps)
    model.train()
    out = {"losses": [], "logits": []}
    for _ in range(steps):
        loss, logits = train_step(model, opt, sched, ids, mask, labels, clip)
        out["losses"].append(loss)
        out["logits"].append(logits.double().cpu())
    if keep_grads:
        out["grads"] = {k: p.grad.detach().double().cpu() for k, p in model.named_parameters() if p.grad is not None}
    model.eval()
    ids2, mask2, _ = make_batch(n, s, SEED_DATA + 1, device)
    out["eval_logits"] = eval_step(model, ids2, mask2).double().cpu()
    model.train()
    return out


WATCH = ("bert.embeddings.word_embeddings.weight", "bert.embeddings.position_embeddings.weight",
         "bert.embeddings.LayerNorm.weight", "bert.encoder.layer.0.attention.self.query.weight",
         "bert.encoder.layer.0.attention.self.value.bias", "bert.encoder.layer.0.attention.output.LayerNorm.bias",
         "bert.encoder.layer.1.intermediate.dense.weight", "bert.encoder.layer.1.output.dense.bias",
         "bert.pooler.dense.weight", "classifier.weight", "classifier.bias")
