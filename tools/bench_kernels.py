"""Per-kernel timings for the layer shapes of PyramidNet (batch 64) -- CUDA events, warm-up,
L2 flushed between iterations.  Prints microseconds and achieved TFLOP/s or GB/s."""
import sys
import torch
sys.path.insert(0, ".")
from distributed_training_dl_b200 import _native as nv
from distributed_training_dl_b200.ops import Conv2d, functional as fn
from distributed_training_dl_b200.utils.timing import flush_l2


def timeit(f, reps=20, flush=True):
    """Captures `reps` back-to-back calls in a CUDA graph (no CPU launch gaps), replays it 3x."""
    for _ in range(2):
        f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            f()
    best = 1e30
    for _ in range(3):
        if flush:
            flush_l2()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best / reps * 1e3  # us per call


def main():
    N = 64
    import os
    prec = os.environ.get("PRECISION", "bf16")
    dt = fn.act_dtype(prec)
    act = fn.act_of(dt)
    esz = 4 if act else 2
    print(f"precision {prec}")
    shapes = eval(os.environ['SHAPES']) if 'SHAPES' in os.environ else [(16, 21, 32), (56, 61, 32), (96, 101, 32), (141, 146, 16), (181, 186, 16), (226, 231, 8), (266, 271, 8)]
    print(f"{'shape':>22s} {'fprop us':>9s} {'TF/s':>6s} {'dgrad us':>9s} {'wgrad us':>9s} {'TF/s':>6s} {'bn fwd':>7s} {'bn bwd':>7s} {'GB/s':>6s}")
    for cin, cout, hw in shapes:
        conv = Conv2d(cin, cout, 3, padding=1, bias=False).cuda()
        conv._ensure_packed(dt)
        x = torch.randn(N, hw, hw, conv.Cin_p, device="cuda").to(dt)
        dy = torch.randn(N, hw, hw, conv.Cout_p, device="cuda").to(dt)
        flops = 2.0 * N * hw * hw * cin * cout * 9
        t_f = timeit(lambda: fn.conv_fprop_raw(x, conv.wf, conv.Cout_p, 3, 3, 1))
        t_d = timeit(lambda: fn.conv_fprop_raw(dy, conv.wd, conv.Cin_p, 3, 3, 1))
        gw = torch.zeros_like(conv.weight)
        need = nv.lib().dtdl_conv_wgrad_workspace_bytes_ex(N, hw, hw, conv.Cin_p, conv.Cout_p, 3, 3, 1, act)
        ws = fn.scratch(x.device, need)
        t_w = timeit(lambda: nv.call("conv_wgrad", nv.ptr(x), nv.ptr(dy), nv.ptr(gw), nv.ptr(ws), ws.numel(), N, hw, hw,
                                     cin, conv.Cin_p, hw, hw, cout, conv.Cout_p, 3, 3, 1, 1, 0, 1, act, nv.stream_ptr()))
        Cp, P = conv.Cout_p, N * hw * hw
        stat = torch.zeros(4 * Cp, device="cuda"); save = torch.zeros(2, Cp, device="cuda")
        gam = torch.ones(cout, device="cuda"); bet = torch.zeros(cout, device="cuda")
        yb = torch.empty_like(dy); dxb = torch.empty_like(dy); gg = torch.zeros(cout, device="cuda"); gb = torch.zeros(cout, device="cuda")

        def bn_fwd():
            nv.call("bn_stats", nv.ptr(dy), nv.ptr(stat), P, Cp, act, nv.stream_ptr())
            nv.call("bn_apply", nv.ptr(dy), nv.ptr(yb), nv.ptr(stat), nv.ptr(gam), nv.ptr(bet), 0, 0, 0, nv.ptr(save[0]),
                    nv.ptr(save[1]), 0, 0, P, cout, Cp, 1e-5, 0.1, 1, 1, stat[2 * Cp:].data_ptr(), 2 * Cp, act, nv.stream_ptr())

        def bn_bwd():
            nv.call("bn_bwd", nv.ptr(x_like), nv.ptr(dy), nv.ptr(yb), nv.ptr(save[0]), nv.ptr(save[1]), nv.ptr(gam),
                    stat[2 * Cp:].data_ptr(), nv.ptr(dxb), 0, 0, nv.ptr(gg), nv.ptr(gb), 0, P, cout, Cp, 1, nv.ptr(stat), 2 * Cp, act, nv.stream_ptr())

        x_like = torch.randn_like(dy)
        t_bf = timeit(bn_fwd)
        t_bb = timeit(bn_bwd)
        if os.environ.get("OVERLAP"):
            # does the weight-gradient kernel overlap with the data-gradient + BN-backward chain when
            # they sit on two branches of the graph?  (serial = one stream, fork = side stream)
            side = torch.cuda.Stream()

            def wgrad_on(stream_ptr):
                nv.call("conv_wgrad", nv.ptr(x), nv.ptr(dy), nv.ptr(gw), nv.ptr(ws), ws.numel(), N, hw, hw, cin,
                        conv.Cin_p, hw, hw, cout, conv.Cout_p, 3, 3, 1, 1, 0, 1, act, stream_ptr)

            def serial():
                wgrad_on(nv.stream_ptr())
                fn.conv_fprop_raw(dy, conv.wd, conv.Cin_p, 3, 3, 1)
                bn_bwd()

            def forked():
                cur = torch.cuda.current_stream()
                side.wait_stream(cur)
                wgrad_on(side.cuda_stream)
                fn.conv_fprop_raw(dy, conv.wd, conv.Cin_p, 3, 3, 1)
                bn_bwd()
                cur.wait_stream(side)

            t_s, t_k = timeit(serial), timeit(forked)
            print(f"{str((cin, cout, hw)):>22s} wgrad+dgrad+bn_bwd: one stream {t_s:7.1f} us, forked {t_k:7.1f} us "
                  f"({100 * (1 - t_k / t_s):.0f} % less)")
        nbytes = N * hw * hw * conv.Cout_p * esz
        print(f"{str((cin, cout, hw)):>22s} {t_f:9.1f} {flops / t_f / 1e6:6.0f} {t_d:9.1f} {t_w:9.1f} {flops / t_w / 1e6:6.0f} "
              f"{t_bf:7.1f} {t_bb:7.1f} {nbytes * 9 / (t_bf + t_bb) / 1e3:6.0f}")


if __name__ == "__main__":
    main()
