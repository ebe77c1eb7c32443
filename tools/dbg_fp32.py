import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, b200net
from oracle import torch_oracle as O
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
b200net.set_precision("fp32")
ref = O.build_reference_model(dropout=False).cuda().train()
ours = b200net.models.resnet50(); ours.fc = O.make_head(); ours.fc[2].p = 0.0
ours.load_state_dict(ref.state_dict()); ours.cuda().train()
x, y = O.make_batch(32, 160, 0, "cuda")
def rel(a, b): return ((a.double() - b.double()).abs().max() / b.double().abs().max()).item()
with torch.no_grad():
    a, b = ours.conv1(x), ref.conv1(x); print("conv1", rel(a, b))
    a2, b2 = ours.bn1(b), ref.bn1(b); print("bn1(same input)", rel(a2, b2))
    a3, b3 = ours.maxpool(torch.relu(b2)), ref.maxpool(torch.relu(b2)); print("maxpool", rel(a3, b3))
    cur = b3
    for li, (lo, lr) in enumerate(zip((ours.layer1, ours.layer2, ours.layer3, ours.layer4), (ref.layer1, ref.layer2, ref.layer3, ref.layer4))):
        for bi, (bo, br) in enumerate(zip(lo, lr)):
            t = br.conv1(cur); print("  L%d.%d conv1" % (li + 1, bi), rel(bo.conv1(cur), t))
            t1 = br.bn1(t); print("  L%d.%d bn1" % (li + 1, bi), rel(bo.bn1(t), t1)); t1 = torch.relu(t1)
            t2 = br.conv2(t1); print("  L%d.%d conv2" % (li + 1, bi), rel(bo.conv2(t1), t2))
            t2b = br.bn2(t2); print("  L%d.%d bn2" % (li + 1, bi), rel(bo.bn2(t2), t2b)); t2b = torch.relu(t2b)
            t3 = br.conv3(t2b); print("  L%d.%d conv3" % (li + 1, bi), rel(bo.conv3(t2b), t3))
            t3b = br.bn3(t3); print("  L%d.%d bn3" % (li + 1, bi), rel(bo.bn3(t3), t3b))
            idt = cur
            if br.downsample is not None:
                d0 = br.downsample[0](cur); print("  L%d.%d ds.conv" % (li + 1, bi), rel(bo.downsample[0](cur), d0))
                idt = br.downsample[1](d0); print("  L%d.%d ds.bn" % (li + 1, bi), rel(bo.downsample[1](d0), idt))
            cur = torch.relu(t3b + idt)
    print("avgpool", rel(ours.avgpool(cur), ref.avgpool(cur)))
    f = torch.flatten(ref.avgpool(cur), 1)
    print("head", rel(b200net.nn.run_head(ours.fc, f), ref.fc(f)))
    print("whole fwd", rel(ours(x), ref(x)))
