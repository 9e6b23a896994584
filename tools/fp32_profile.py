import sys; sys.path.insert(0,'.')
import torch, json
from artspeech_b200 import models, synthetic
torch.manual_seed(0)
net = models.ArtSpeech(45,10,precision="fp32").cuda().eval()
x = synthetic.tokens(256,100,45,seed=0).cuda(); L=[100]*256
ld = torch.tensor(L, dtype=torch.int32, device="cuda")
with torch.no_grad():
    for _ in range(2): net(x, L, lengths_dev=ld)
    net.profile(True); net.read_profile(reset=True)
    for _ in range(5):
        torch.cuda._sleep(400000); net(x, L, lengths_dev=ld)
    torch.cuda.synchronize()
    p = net.read_profile(reset=True)
print(json.dumps({k: v[0]/v[1] for k,v in p.items()}))
