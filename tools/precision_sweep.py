"""PSNR / EPE of tensor-core precision modes against the fp32 CUDA-core path over a clip
(recurrence included).  Usage: precision_sweep.py H W nframes"""
import os, sys, subprocess, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

def run(mode, H, W, n):
    import torch
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Finetunemodel, Network
    sd = synth.synthetic_state_dict(Network(synth.make_args()).state_dict(), seed=1234)
    ck = "/tmp/zt_prec_%d.pt" % os.getpid()
    torch.save(sd, ck)
    m = Finetunemodel(synth.make_args(of_scale=3, model_pretrain=ck)).cuda().eval()
    m.raft.load_state_dict({k[5:]: v for k, v in sd.items() if k.startswith("raft.")})
    clip = synth.synthetic_clip(n, H, W, seed=5).cuda()
    outs = []
    for t in range(n):
        m.is_new_seq = t == 0
        H2, H3, s3 = m(clip[t:t + 1])
        fu = [v for k, v in m.raft.engine().ws.bufs.items() if k[0] == "flow_up"] if t > 0 else []
        outs.append((H2.cpu(), H3.cpu(), fu[0].cpu().clone() if fu else None))
    torch.save(outs, "/tmp/zt_prec_%s.pt" % mode)

if __name__ == "__main__":
    if len(sys.argv) > 4:
        run(sys.argv[4], int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]))
        sys.exit(0)
    import torch
    H, W, n = sys.argv[1:4]
    modes = {"ref": dict(RAFT=0, D1=0, ENH=0, D2=0), "nets1": dict(RAFT=3, D1=1, ENH=1, D2=1),
             "upd1": dict(RAFT=3, RAFT_UPD=1, D1=1, ENH=1, D2=1), "enc1": dict(RAFT=3, RAFT_ENC=1, D1=1, ENH=1, D2=1),
             "head1": dict(RAFT=3, RAFT_HEAD=1, D1=1, ENH=1, D2=1),
             "upd_head1": dict(RAFT=3, RAFT_UPD=1, RAFT_HEAD=1, D1=1, ENH=1, D2=1)}
    for name, env in modes.items():
        e = dict(os.environ); e.update({"ZT_TC_" + k: str(v) for k, v in env.items()})
        subprocess.run([sys.executable, __file__, H, W, n, name], env=e, check=True)
    ref = torch.load("/tmp/zt_prec_ref.pt")
    def psnr(a, b):
        mse = ((a.double() - b.double()) ** 2).mean().item()
        return 200.0 if mse == 0 else 10 * torch.log10(torch.tensor(1.0 / mse)).item()
    for name in modes:
        if name == "ref": continue
        o = torch.load("/tmp/zt_prec_%s.pt" % name)
        p2 = [psnr(a[0], b[0]) for a, b in zip(o, ref)]
        p3 = [psnr(a[1], b[1]) for a, b in zip(o, ref)]
        ep = [((a[2] - b[2]) ** 2).sum(1).sqrt().mean().item() for a, b in zip(o, ref) if a[2] is not None]
        print("%-9s H2 min %.1f dB (last %.1f)  H3 min %.1f dB (last %.1f)  EPE max %.5f mean %.5f" % (
            name, min(p2), p2[-1], min(p3), p3[-1], max(ep), sum(ep) / len(ep)), flush=True)
