#!/usr/bin/env python
"""The reference's image script with ONE import swapped: `from b200net import models` for `from torchvision import
models`.  Everything else is the reference's own statements -- loaders (another_neural_net.py:37-61,
pytorch_training_inference_on_image.ipynb:324-349), freeze + head + NLLLoss + Adam (:105-114 / ipynb:436-446), train
loop with eval (:117-160 / ipynb:657-698), torch.save / torch.load (ipynb:700, :852), predict_image /
get_random_images timing loop (:180-217 / ipynb:855-905) -- with exactly the API-drift shims SURVEY.md 8(c) lists
(`next(dataiter)` for `dataiter.next()`, `--local-rank` as torch.distributed.launch now spells it) and sizes / epochs
exposed as flags so the demo finishes in seconds.

    python -m torch.distributed.launch --nproc_per_node=N scripts/image_train_infer.py --data DIR
    python scripts/image_train_infer.py --synthetic          # writes a small Imagenette-shaped folder of JPEGs first

Flags beyond the reference: --ddp wraps the model in the DistributedDataParallel constructor the reference leaves
commented out (pytorch_on_language_distr.py:220-221); --all-layers trains every layer with SGD(lr=0.001), the
resnet.py:17-25 recipe, instead of the frozen-backbone recipe."""
import argparse
import os
import sys
import tempfile
import time

import numpy as np
import torch
from torch import nn
from torch import optim
from torch.autograd import Variable
from torch.utils.data.distributed import DistributedSampler
from torchvision import datasets, transforms

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200net import models                      # was: from torchvision import models
import b200net

parser = argparse.ArgumentParser()
parser.add_argument("--local_rank", "--local-rank", default=int(os.environ.get("LOCAL_RANK", 0)), type=int)
parser.add_argument("--data", default=None, help="ImageFolder root (imagenette2/train in the reference)")
parser.add_argument("--synthetic", action="store_true", help="write a small synthetic image folder and use it")
parser.add_argument("--size", default=224, type=int)          # Resize((224,224)) in the reference
parser.add_argument("--batch", default=64, type=int)          # batch_size=64 in the reference
parser.add_argument("--epochs", default=1, type=int)
parser.add_argument("--infer-images", default=1000, type=int)  # get_random_images(1000)
parser.add_argument("--ddp", action="store_true")
parser.add_argument("--all-layers", action="store_true")
args = parser.parse_args()

distributed = "RANK" in os.environ and "WORLD_SIZE" in os.environ
if distributed:
    # the reference: torch.distributed.init_process_group('gloo'); NCCL is the intended backend on GPUs
    torch.distributed.init_process_group("nccl" if torch.cuda.is_available() else "gloo")


def make_synthetic_folder(root, classes=10, per_class=12, seed=0):
    from PIL import Image
    rng = np.random.RandomState(seed)
    for c in range(classes):
        d = os.path.join(root, "n%08d" % c)
        os.makedirs(d, exist_ok=True)
        for i in range(per_class):
            h, w = rng.randint(160, 400), rng.randint(160, 500)
            img = (rng.rand(h, w, 3) * 255 * (0.3 + 0.7 * c / classes)).astype(np.uint8)
            Image.fromarray(img).save(os.path.join(d, "img_%03d.JPEG" % i), quality=85)
    return root


data_dir = args.data
if args.synthetic or data_dir is None:
    data_dir = os.path.join(tempfile.gettempdir(), "b200net_synthetic_imagenette")
    if args.local_rank == 0 and not os.path.isdir(data_dir):
        make_synthetic_folder(data_dir)
    if distributed:
        torch.distributed.barrier()


def load_split_train_test(datadir, valid_size=.2):
    train_transforms = transforms.Compose([transforms.Resize((args.size, args.size)), transforms.ToTensor(), ])
    test_transforms = transforms.Compose([transforms.Resize((args.size, args.size)), transforms.ToTensor(), ])
    train_data = datasets.ImageFolder(datadir, transform=train_transforms)
    test_data = datasets.ImageFolder(datadir, transform=test_transforms)
    num_train = len(train_data)
    indices = list(range(num_train))
    split = int(np.floor(valid_size * num_train))
    np.random.shuffle(indices)
    from torch.utils.data.sampler import SubsetRandomSampler
    train_idx, test_idx = indices[split:], indices[:split]
    if distributed:
        # another_neural_net.py:54-55 (the sampler walks positions of the index list; SURVEY 0.2 notes the overlap)
        train_sampler = DistributedSampler(train_idx)
        test_sampler = DistributedSampler(test_idx)
    else:
        train_sampler = SubsetRandomSampler(train_idx)      # ...on_image.ipynb:341-342
        test_sampler = SubsetRandomSampler(test_idx)
    trainloader = torch.utils.data.DataLoader(train_data, sampler=train_sampler, batch_size=args.batch)
    testloader = torch.utils.data.DataLoader(test_data, sampler=test_sampler, batch_size=args.batch)
    return trainloader, testloader


trainloader, testloader = load_split_train_test(data_dir, .2)
device = torch.device("cuda", args.local_rank)
torch.cuda.set_device(device)

model = models.resnet50(pretrained=False)       # pretrained=True in the reference: needs the cached checkpoint file
if not args.all_layers:
    for param in model.parameters():
        param.requires_grad = False
model.fc = nn.Sequential(nn.Linear(2048, 512),
                         nn.ReLU(),
                         nn.Dropout(0.2),
                         nn.Linear(512, 10),
                         nn.LogSoftmax(dim=1))
criterion = nn.NLLLoss()
if args.all_layers:
    optimizer = b200net.optim.SGD(model.parameters(), lr=0.001)             # resnet.py:24
else:
    optimizer = optim.Adam(model.fc.parameters(), lr=0.003)
model.to(device)
net = model
if args.ddp and distributed:
    net = b200net.parallel.DistributedDataParallel(model, device_ids=[args.local_rank], output_device=args.local_rank)

t1 = time.time()
epochs = args.epochs
steps = 0
running_loss = 0
print_every = 10
train_losses, test_losses = [], []
for epoch in range(epochs):
    for inputs, labels in trainloader:
        steps += 1
        inputs, labels = inputs.to(device), labels.to(device)
        optimizer.zero_grad()
        logps = net.forward(inputs)
        loss = criterion(logps, labels)
        loss.backward()
        optimizer.step()
        running_loss += loss.item()
    test_loss = 0
    accuracy = 0
    net.eval()
    with torch.no_grad():
        for inputs, labels in testloader:
            inputs, labels = inputs.to(device), labels.to(device)
            logps = net.forward(inputs)
            batch_loss = criterion(logps, labels)
            test_loss += batch_loss.item()
            ps = torch.exp(logps)
            top_p, top_class = ps.topk(1, dim=1)
            equals = top_class == labels.view(*top_class.shape)
            accuracy += torch.mean(equals.type(torch.FloatTensor)).item()
    train_losses.append(running_loss / len(trainloader))
    test_losses.append(test_loss / len(testloader))
    print(f"Epoch {epoch + 1}/{epochs}.. "
          f"Train loss: {running_loss / print_every:.3f}.. "
          f"Test loss: {test_loss / len(testloader):.3f}.. "
          f"Test accuracy: {accuracy / len(testloader):.3f}")
    running_loss = 0
    net.train()
ckpt = os.path.join(tempfile.gettempdir(), "b200net_aerialmodel_rank%d.pth" % args.local_rank)
torch.save(model, ckpt)
print("Time to train is {} seconds".format(time.time() - t1))

test_transforms = transforms.Compose([transforms.Resize((args.size, args.size)), transforms.ToTensor()])
model = torch.load(ckpt, weights_only=False)
model.eval()
model = model.to(device)


def predict_image(image):
    image_tensor = test_transforms(image).float()
    image_tensor = image_tensor.unsqueeze_(0)
    input = Variable(image_tensor)
    input = input.to(device)
    output = model(input)
    index = output.data.cpu().numpy().argmax()
    return index


def get_random_images(num):
    data = datasets.ImageFolder(data_dir, transform=test_transforms)
    indices = list(range(len(data)))
    np.random.shuffle(indices)
    idx = indices[:num]
    from torch.utils.data.sampler import SubsetRandomSampler
    sampler = SubsetRandomSampler(idx)
    loader = torch.utils.data.DataLoader(data, sampler=sampler, batch_size=num)
    dataiter = iter(loader)
    images, labels = next(dataiter)                   # was: dataiter.next()
    return images, labels


t1 = time.time()
to_pil = transforms.ToPILImage()
images, labels = get_random_images(args.infer_images)
correct = 0
for ii in range(len(images)):
    image = to_pil(images[ii])
    index = predict_image(image)
    res = int(labels[ii]) == index
    correct += int(res)
print("Time to infer %d images is %s seconds" % (len(images), time.time() - t1))
if distributed:
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()
print("image_train_infer: OK")
