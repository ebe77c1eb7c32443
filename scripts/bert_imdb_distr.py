#!/usr/bin/env python
"""pytorch_on_language_distr.py with the model / optimizer imports swapped for b200net's and the commented-out
DistributedDataParallel line ENABLED (:220-221) -- everything else is the reference's own statements: --local_rank
(:21-23), init_process_group (:26), rm_tags + CSV read (:34-55), tokenise / pad to MAX_LEN / attention masks
(:56-103), train / validation split (:105-112), TensorDataset + DistributedSampler + DataLoader (:134-149),
BertForSequenceClassification.from_pretrained (:155-161), AdamW(lr=2e-5, eps=1e-8) + linear schedule (:167-183),
seeds (:212-217), the training loop with clip_grad_norm_ (:244-275), validation (:296-335) and the timed test loop
(:342-379).  Shims: keras' pad_sequences -> b200net.text.pad_sequences; transformers.AdamW (gone
in transformers 5) -> b200net.bert.AdamW; the tokenizer vocabulary and the IMDB csv cannot be downloaded here, so
without --csv a synthetic corpus is generated and the WordPiece vocabulary is built from it.

    python -m torch.distributed.launch --nproc_per_node=N scripts/bert_imdb_distr.py [--csv "IMDB Dataset.csv"]
    python scripts/bert_imdb_distr.py --layers 2 --reviews 256        # seconds-long demo on one GPU"""
import argparse
import datetime
import os
import random
import sys
import tempfile
import time

import numpy as np
import pandas as pd
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200net
from b200net.bert import BertForSequenceClassification, AdamW, get_linear_schedule_with_warmup   # was: transformers
from b200net.text import rm_tags, pad_sequences, build_wordpiece_vocab                          # was: keras / inline

parser = argparse.ArgumentParser()
parser.add_argument("--local_rank", "--local-rank", type=int, default=int(os.environ.get("LOCAL_RANK", 0)))
parser.add_argument("--csv", default=None, help='"IMDB Dataset.csv" (review, sentiment)')
parser.add_argument("--reviews", type=int, default=512, help="synthetic corpus size when no csv is given")
parser.add_argument("--layers", type=int, default=12, help="encoder layers (12 = bert-base)")
parser.add_argument("--epochs", type=int, default=1)          # 3 in the reference
parser.add_argument("--max-len", type=int, default=128)       # MAX_LEN = 128
parser.add_argument("--pretrained", default=None, help="local Hugging Face checkpoint directory")
args = parser.parse_args()

distributed = "RANK" in os.environ and "WORLD_SIZE" in os.environ
if distributed:
    torch.distributed.init_process_group("nccl")               # 'gloo' in the reference
device = torch.device("cuda", args.local_rank)
torch.cuda.set_device(device)


def synthetic_reviews(n, seed=0):
    rng = random.Random(seed)
    pos = "great wonderful moving brilliant superb loved touching masterpiece".split()
    neg = "boring awful dull terrible waste hated clumsy mess".split()
    filler = ("the film movie plot actor scene story was is and a of in it this that with for camera script ending "
              "director character music").split()
    rows = []
    for _ in range(n):
        label = rng.random() < 0.5
        words = [rng.choice(filler) if rng.random() < 0.8 else rng.choice(pos if label else neg)
                 for _ in range(rng.randint(20, 200))]
        rows.append((" ".join(words[:10]) + " <br /><br /> " + " ".join(words[10:]), "positive" if label else "negative"))
    return pd.DataFrame(rows, columns=["review", "sentiment"])


df = pd.read_csv(args.csv) if args.csv else synthetic_reviews(args.reviews)
df["review"] = df["review"].apply(rm_tags)
sentences = df.review.values
labels = (df.sentiment.values == "positive").astype(np.int64)

from transformers import BertTokenizer
if args.pretrained and os.path.isfile(os.path.join(args.pretrained, "vocab.txt")):
    tokenizer = BertTokenizer.from_pretrained(args.pretrained, do_lower_case=True)
else:
    vocab = os.path.join(tempfile.gettempdir(), "b200net_vocab_rank%d.txt" % args.local_rank)
    vocab = build_wordpiece_vocab(sentences, vocab)
    try:
        tokenizer = BertTokenizer(vocab=vocab, do_lower_case=True)             # transformers >= 5
    except TypeError:
        tokenizer = BertTokenizer(vocab_file=vocab, do_lower_case=True)        # transformers 4.x

MAX_LEN = args.max_len
input_ids = [tokenizer.encode(sent, add_special_tokens=True, max_length=MAX_LEN, truncation=True) for sent in sentences]
input_ids = pad_sequences(input_ids, maxlen=MAX_LEN, dtype="long", value=0, truncating="post", padding="post")
attention_masks = [[int(token_id > 0) for token_id in sent] for sent in input_ids]

from sklearn.model_selection import train_test_split
train_inputs, validation_inputs, train_labels, validation_labels = train_test_split(input_ids, labels,
                                                                                    random_state=2020, test_size=0.1)
train_masks, validation_masks, _, _ = train_test_split(attention_masks, labels, random_state=2020, test_size=0.1)
train_inputs, validation_inputs = torch.tensor(train_inputs), torch.tensor(validation_inputs)
train_labels, validation_labels = torch.tensor(train_labels), torch.tensor(validation_labels)
train_masks, validation_masks = torch.tensor(train_masks), torch.tensor(validation_masks)

from torch.utils.data import TensorDataset, DataLoader, RandomSampler, SequentialSampler, DistributedSampler
batch_size = 32
train_data = TensorDataset(train_inputs, train_masks, train_labels)
train_sampler = DistributedSampler(train_data) if distributed else RandomSampler(train_data)
train_dataloader = DataLoader(train_data, sampler=train_sampler, batch_size=batch_size)
validation_data = TensorDataset(validation_inputs, validation_masks, validation_labels)
validation_sampler = DistributedSampler(validation_data) if distributed else SequentialSampler(validation_data)
validation_dataloader = DataLoader(validation_data, sampler=validation_sampler, batch_size=batch_size)

if args.pretrained:
    model = BertForSequenceClassification.from_pretrained(args.pretrained, num_labels=2, output_attentions=False,
                                                          output_hidden_states=False)
else:
    model = BertForSequenceClassification(b200net.bert.BertConfig(num_labels=2, num_hidden_layers=args.layers,
                                                                  vocab_size=max(len(tokenizer), 8)))
optimizer = AdamW(model.parameters(), lr=2e-5, eps=1e-8)
epochs = args.epochs
total_steps = len(train_dataloader) * epochs
scheduler = get_linear_schedule_with_warmup(optimizer, num_warmup_steps=0, num_training_steps=total_steps)


def flat_accuracy(preds, labels):
    pred_flat = np.argmax(preds, axis=1).flatten()
    labels_flat = labels.flatten()
    return np.sum(pred_flat == labels_flat) / len(labels_flat)


def format_time(elapsed):
    return str(datetime.timedelta(seconds=int(round((elapsed)))))


seed_val = 42
random.seed(seed_val)
np.random.seed(seed_val)
torch.manual_seed(seed_val)
torch.cuda.manual_seed_all(seed_val)

model = model.to(device)
if distributed:
    # the line the reference leaves commented out (:220-221)
    model = b200net.parallel.DistributedDataParallel(model, device_ids=[args.local_rank], output_device=args.local_rank)

loss_values = []
for epoch_i in range(0, epochs):
    print('======== Epoch {:} / {:} ========'.format(epoch_i + 1, epochs))
    t0 = time.time()
    total_loss = 0
    model.train()
    for step, batch in enumerate(train_dataloader):
        b_input_ids = batch[0].to(device)
        b_input_mask = batch[1].to(device)
        b_labels = batch[2].to(device)
        model.zero_grad()
        outputs = model(b_input_ids, token_type_ids=None, attention_mask=b_input_mask, labels=b_labels)
        loss = outputs[0]
        total_loss += loss.item()
        loss.backward()
        b200net.utils.clip_grad_norm_(model.parameters(), 1.0)      # torch.nn.utils.clip_grad_norm_ in the reference
        optimizer.step()
        scheduler.step()
    avg_train_loss = total_loss / len(train_dataloader)
    loss_values.append(avg_train_loss)
    print("  Average training loss: {0:.2f}".format(avg_train_loss))
    print("  Training epcoh took: {:}".format(format_time(time.time() - t0)))

    t0 = time.time()
    model.eval()
    eval_loss, eval_accuracy = 0, 0
    nb_eval_steps, nb_eval_examples = 0, 0
    for batch in validation_dataloader:
        batch = tuple(t.to(device) for t in batch)
        b_input_ids, b_input_mask, b_labels = batch
        with torch.no_grad():
            outputs = model(b_input_ids, token_type_ids=None, attention_mask=b_input_mask)
        logits = outputs[0]
        logits = logits.detach().cpu().numpy()
        label_ids = b_labels.to('cpu').numpy()
        eval_accuracy += flat_accuracy(logits, label_ids)
        nb_eval_steps += 1
    print("  Accuracy: {0:.2f}".format(eval_accuracy / max(1, nb_eval_steps)))
    print("  Validation took: {:}".format(format_time(time.time() - t0)))

print("Training complete!")
if distributed:
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()
print("bert_imdb_distr: OK")
