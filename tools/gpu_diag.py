"""GPU diagnostic: CUDA path vs C oracle vs reference CUDA build, with timings.  Development aid."""
import os, sys, time, traceback
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
import gsr_synth as S
import diff_gaussian_rasterization as D
from oracle import oracle as O, ref_cuda as RC

dev = torch.device("cuda:0")

def settings(cam, bg, deg, debug=False):
    return D.GaussianRasterizationSettings(image_height=cam["H"], image_width=cam["W"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
        bg=bg.to(dev), scale_modifier=1.0, viewmatrix=cam["viewmatrix"].to(dev), projmatrix=cam["projmatrix"].to(dev), sh_degree=deg,
        campos=cam["campos"].to(dev), prefiltered=False, debug=debug)

def rel_err(a, b):
    a = a.double().reshape(-1); b = b.double().reshape(-1)
    d = (a - b).abs()
    return float(d.max()), float(d.max() / (b.abs().max() + 1e-30)), float((d / (b.abs() + 1e-6 * b.abs().max() + 1e-30)).median())

def run_case(P, W, H, seed, cam, bg, dense_ct=True, s0=None):
    print(f"=== case P={P} {W}x{H} seed={seed} bg={bg.tolist()} dense_ct={dense_ct}", flush=True)
    sc = S.make_gvol(P, W, H, seed=seed, s0=s0)
    ct = S.cotangents(W, H, seed=seed + 100, dense=dense_ct)
    leaves = {k: sc[k].to(dev).clone().requires_grad_(True) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
    m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
    rs = settings(cam, bg, 3, debug=True)
    rast = D.GaussianRasterizer(rs)
    color, radii, depth, alpha = rast(leaves["means3D"], m2d, leaves["opacities"], shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"])
    torch.cuda.synchronize()
    fn = color.grad_fn
    saved = fn.saved_tensors
    geom, binning, img = saved[7], saved[8], saved[9]
    R = fn.num_rendered
    mine = D.rasterizer_state(P, W, H, R, geom, binning, img)
    print("mine: R", R, "visible", int((radii > 0).sum()))
    # ---- oracle
    st = O.forward(sc["means3D"].numpy(), sc["opacities"].numpy(), shs=sc["shs"].numpy(), scales=sc["scales"].numpy(), rotations=sc["rotations"].numpy(),
                   viewmatrix=cam["viewmatrix"].numpy(), projmatrix=cam["projmatrix"].numpy(), campos=cam["campos"].numpy(), bg=bg.numpy(), W=W, H=H,
                   tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], sh_degree=3)
    print("oracle: R", st["num_rendered"])
    print(" radii mismatches vs oracle:", int((radii.cpu().numpy() != st["radii"]).sum()))
    vis = st["radii"] > 0
    print(" depth bits mismatches:", int((mine["depths"].cpu().numpy().view(np.uint32)[vis] != st["depths"].view(np.uint32)[vis]).sum()),
          " xy bits mismatches:", int((mine["means2D"].cpu().numpy().view(np.uint32)[vis] != st["xy"].view(np.uint32)[vis]).sum()),
          " conic bits mismatches:", int((mine["conic_opacity"].cpu().numpy().view(np.uint32)[vis] != st["conic_opacity"].view(np.uint32)[vis]).sum()),
          " tiles mismatches:", int((mine["tiles_touched"].cpu().numpy().astype(np.uint32) != st["tiles_touched"]).sum()))
    print(" rgb maxdiff:", float(np.abs(mine["rgb"].cpu().numpy()[vis] - st["rgb"][vis]).max()))
    if R == st["num_rendered"]:
        print(" keys equal:", bool(np.array_equal(mine["keys_sorted"].cpu().numpy().view(np.uint64), st["keys_sorted"])),
              " point_list equal:", bool(np.array_equal(mine["point_list"].cpu().numpy().view(np.uint32), st["point_list"])),
              " ranges equal:", bool(np.array_equal(mine["ranges"].cpu().numpy().view(np.uint32), st["ranges"])))
    for name, a, b in [("color", color, st["color"]), ("depth", depth, st["depth"]), ("alpha", alpha, st["alpha"])]:
        d = np.abs(a.detach().cpu().numpy() - b)
        print(f" {name}: maxdiff {d.max():.3e}  pixels>1e-5: {int((d > 1e-5).sum())}")
    print(" n_contrib mismatches:", int((mine["n_contrib"].cpu().numpy().view(np.uint32) != st["n_contrib"]).sum()))
    gd = ct["grad_depth"].to(dev) if dense_ct else None
    ga = ct["grad_alpha"].to(dev) if dense_ct else None
    loss = (color * ct["grad_color"].to(dev)).sum()
    if dense_ct:
        loss = loss + (depth * gd).sum() + (alpha * ga).sum()
    loss.backward()
    torch.cuda.synchronize()
    go = O.backward(st, ct["grad_color"].numpy(), ct["grad_depth"].numpy(), ct["grad_alpha"].numpy())
    for name, key in [("means3D", "means3D"), ("opacities", "opacities"), ("shs", "sh"), ("scales", "scales"), ("rotations", "rotations")]:
        print(f" grad {name} vs oracle: max abs %.3e  rel-to-max %.3e  median rel %.3e" % rel_err(leaves[name].grad.cpu(), torch.from_numpy(go[key])))
    print(" grad means2D vs oracle: max abs %.3e  rel-to-max %.3e  median rel %.3e" % rel_err(m2d.grad.cpu(), torch.from_numpy(go["means2D"])))
    # ---- reference CUDA
    if RC.available():
        try:
            kw = dict(viewmatrix=cam["viewmatrix"].to(dev), projmatrix=cam["projmatrix"].to(dev), campos=cam["campos"].to(dev), bg=bg.to(dev), W=W, H=H,
                      tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], sh_degree=3)
            rst = RC.forward(sc["means3D"].to(dev), sc["opacities"].to(dev), shs=sc["shs"].to(dev), scales=sc["scales"].to(dev), rotations=sc["rotations"].to(dev), **kw)
            torch.cuda.synchronize()
            ref = RC.parse_state(rst, P, W, H)
            print("reference CUDA: R", rst["num_rendered"])
            print(" radii mismatches vs ref:", int((radii != rst["radii"]).sum()))
            if R == rst["num_rendered"]:
                print(" keys equal:", bool(torch.equal(mine["keys_sorted"], ref["keys_sorted"])), " point_list equal:", bool(torch.equal(mine["point_list"], ref["point_list"])),
                      " ranges equal:", bool(torch.equal(mine["ranges"], ref["ranges"])))
            for name, a, b in [("color", color, rst["color"]), ("depth", depth, rst["depth"]), ("alpha", alpha, rst["alpha"])]:
                d = (a.detach() - b).abs()
                print(f" {name}: maxdiff {float(d.max()):.3e}  pixels>1e-5: {int((d > 1e-5).sum())}  bit-equal: {bool(torch.equal(a.detach(), b))}")
            print(" n_contrib mismatches:", int((mine["n_contrib"] != ref["n_contrib"]).sum()))
            visg = radii > 0
            print(" rgb maxdiff vs ref:", float((mine["rgb"][visg] - ref["rgb"][visg]).abs().max()), " conic bit mismatches:",
                  int((mine["conic_opacity"][visg].view(torch.int32) != ref["conic_opacity"][visg].view(torch.int32)).sum()))
            gr = RC.backward(rst, ct["grad_color"].to(dev), ct["grad_depth"].to(dev), ct["grad_alpha"].to(dev))
            torch.cuda.synchronize()
            for name, key in [("means3D", "means3D"), ("opacities", "opacities"), ("shs", "sh"), ("scales", "scales"), ("rotations", "rotations")]:
                print(f" grad {name} vs ref: max abs %.3e  rel-to-max %.3e  median rel %.3e" % rel_err(leaves[name].grad, gr[key]))
            print(" grad means2D vs ref: max abs %.3e  rel-to-max %.3e  median rel %.3e" % rel_err(m2d.grad, gr["means2D"]))
            for name, key in [("means3D", "means3D"), ("opacities", "opacities"), ("shs", "sh")]:
                print(f" [oracle vs ref] grad {name}: max abs %.3e  rel-to-max %.3e  median rel %.3e" % rel_err(torch.from_numpy(go[key]).to(dev), gr[key]))
        except Exception:
            traceback.print_exc()
    else:
        print("reference CUDA build not available")

def timing(P, W, H, seed):
    print(f"=== timing P={P} {W}x{H}", flush=True)
    cam = S.make_camera(W, H)
    sc = S.make_gvol(P, W, H, seed=seed)
    bg = torch.zeros(3)
    ct = S.cotangents(W, H, seed=1, dense=False)
    gc = ct["grad_color"].to(dev)
    t = {k: sc[k].to(dev) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
    rs = settings(cam, bg, 3)
    def mine_step():
        leaves = {k: v.detach().requires_grad_(True) for k, v in t.items()}
        m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
        color, radii, depth, alpha = D.GaussianRasterizer(rs)(leaves["means3D"], m2d, leaves["opacities"], shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"])
        color.backward(gc)
    def mine_fwd():
        with torch.no_grad():
            D.GaussianRasterizer(rs)(t["means3D"], None, t["opacities"], shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    kw = dict(viewmatrix=cam["viewmatrix"].to(dev), projmatrix=cam["projmatrix"].to(dev), campos=cam["campos"].to(dev), bg=bg.to(dev), W=W, H=H,
              tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], sh_degree=3)
    zeros1 = torch.zeros(1, H, W, device=dev)
    def ref_fwd():
        return RC.forward(t["means3D"], t["opacities"], shs=t["shs"], scales=t["scales"], rotations=t["rotations"], **kw)
    def ref_step():
        st = ref_fwd()
        RC.backward(st, gc, zeros1, zeros1)
    def bench(fn, n=10):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n): fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    print(" mine fwd ms %.3f   fwd+bwd ms %.3f" % (bench(mine_fwd), bench(mine_step)), flush=True)
    if RC.available():
        print(" ref  fwd ms %.3f   fwd+bwd ms %.3f" % (bench(ref_fwd), bench(ref_step)), flush=True)

if __name__ == "__main__":
    print(torch.cuda.get_device_name(0), "lib version", D.lib.gsr_version(), "ref available", RC.available(), flush=True)
    try:
        run_case(20000, 320, 180, 11, S.make_camera(320, 180), torch.tensor([0., 0., 0.]))
        run_case(20000, 320, 180, 12, S.second_view(320, 180), torch.tensor([1., 1., 1.]), s0=0.03)
        run_case(100000, 640, 360, 1, S.make_camera(640, 360), torch.tensor([0., 0., 0.]), dense_ct=False)
    except Exception:
        traceback.print_exc()
    try:
        timing(100000, 640, 360, 1)
        timing(1000000, 1920, 1080, 5)
    except Exception:
        traceback.print_exc()
