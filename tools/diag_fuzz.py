import sys, os, numpy as np, torch
ROOT="/root/repo"; sys.path.insert(0, ROOT+"/cf-3dgs_b200"); sys.path.insert(0, ROOT); sys.path.insert(0, ROOT+"/tests")
import gsr_synth as S
from helpers import cuda_forward
from test_gpu_fuzz import _case
from oracle import ref_cuda as RC
i=int(sys.argv[1]) if len(sys.argv)>1 else 0
cam, sc, deg, bg, W, H, P = _case(i)
out, leaves, m2d = cuda_forward(sc, cam, bg, sh_degree=deg)
dev=out["color"].device
rst = RC.forward(sc["means3D"].to(dev), sc["opacities"].to(dev), shs=sc["shs"].to(dev), scales=sc["scales"].to(dev), rotations=sc["rotations"].to(dev), viewmatrix=cam["viewmatrix"].to(dev), projmatrix=cam["projmatrix"].to(dev), campos=cam["campos"].to(dev), bg=torch.tensor(bg, dtype=torch.float32, device=dev), W=W, H=H, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], sh_degree=deg)
ct = S.cotangents(W, H, seed=4000 + i, dense=bool(i % 3))
zero = torch.zeros(1, H, W, device=dev)
gd = ct["grad_depth"].to(dev) if i % 3 else zero; ga = ct["grad_alpha"].to(dev) if i % 3 else zero
loss = (out["color"] * ct["grad_color"].to(dev)).sum()
if i % 3: loss = loss + (out["depth"] * gd).sum() + (out["alpha"] * ga).sum()
loss.backward()
gr = RC.backward(rst, ct["grad_color"].to(dev), gd, ga)
print("case", i, "P", P, W, H, "deg", deg, "R", out["num_rendered"])
for nm,key in [("means3D","means3D"),("scales","scales"),("rotations","rotations"),("opacities","opacities"),("shs","sh")]:
    a=leaves[nm].grad.reshape(P,-1).double().cpu(); b=gr[key].reshape(P,-1).double().cpu()
    d=(a-b).abs().max(1).values
    j=int(d.argmax())
    print(nm, "max|ref|", float(b.abs().max()), "worst idx", j, "err", float(d[j]), "ours", a[j][:4].tolist(), "ref", b[j][:4].tolist())
j=int((leaves["means3D"].grad.double().cpu()-gr["means3D"].double().cpu()).abs().max(1).values.argmax())
print("gaussian", j, "mean", sc["means3D"][j].tolist(), "scale", sc["scales"][j].tolist(), "rot", sc["rotations"][j].tolist(), "opac", sc["opacities"][j].tolist(), "radius", int(out["radii"][j]))
a2=m2d.grad.cpu(); b2=gr["means2D"].cpu(); print("means2D grad ours", a2[j].tolist(), "ref", b2[j].tolist())
st=out["state"]; print("conic_opacity", st["conic_opacity"][j].tolist(), "xy", st["means2D"][j].tolist(), "depth", float(st["depths"][j]))
# ground truth by float64 autograd through the PyTorch restatement
from oracle import torch_oracle as TO
lv={k: sc[k].double().clone().requires_grad_(True) for k in ["means3D","opacities","shs","scales","rotations"]}
o=TO.forward(lv["means3D"], lv["opacities"], shs=lv["shs"], scales=lv["scales"], rotations=lv["rotations"], viewmatrix=cam["viewmatrix"].double(), projmatrix=cam["projmatrix"].double(), campos=cam["campos"].double(), bg=torch.tensor(bg).double(), W=W,H=H,tanfovx=cam["tanfovx"],tanfovy=cam["tanfovy"],sh_degree=deg)
l=(o["color"]*ct["grad_color"].double()).sum()
if i%3: l=l+(o["depth"]*ct["grad_depth"].double()).sum()+(o["alpha"]*ct["grad_alpha"].double()).sum()
l.backward()
print("float64 autograd means3D grad", lv["means3D"].grad[j].tolist())
print("ours", leaves["means3D"].grad[j].tolist(), "ref", gr["means3D"][j].tolist())
