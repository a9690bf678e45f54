import sys, torch
sys.path.insert(0,".")
from poisson_ellipse_openmp_mpi_cuda_b200.ops.gpu import GpuSolver
M,N=200,300
s=GpuSolver(M,N).finalize()
s.step("assemble"); torch.cuda.synchronize()
co=s.col_off
a,b,D=s.array("a"),s.array("b"),s.array("D")
h=max(2.0/M,1.2/N); eps=h*h; ie=1.0/eps
def cls_of(x): return torch.where(x==1.0, 0, torch.where(x==ie, 1, 2))
d_in=2.0/((2.0/M)**2)+2.0/((1.2/N)**2); d_out=(2*ie)/((2.0/M)**2)+(2*ie)/((1.2/N)**2)
kd=torch.where(D==d_in,0,torch.where(D==d_out,1,2))
exp=(cls_of(a)|(cls_of(b)<<2)|(kd<<4)).to(torch.uint8)
for (x,y) in [(0,0),(64,32),(128,0)]:
    st, box = s.tma_probe(12, x, y, 10, 34)   # 10 doubles = 80 bytes per row
    got = box.view(torch.uint8).reshape(34,80)
    ref = torch.zeros_like(got)
    sub = exp[y:y+34, x:x+80]
    ref[:sub.shape[0], :sub.shape[1]] = sub
    print("probe",x,y,"status",st,"match",bool(torch.equal(got,ref)), "nonzero", int((got!=0).sum()), "kd2 frac", float((kd==2).double().mean()), flush=True)
    if not torch.equal(got,ref):
        bad=(got!=ref).nonzero()[:5]; print(bad.tolist(), got[bad[0][0], :16].tolist(), ref[bad[0][0], :16].tolist())
