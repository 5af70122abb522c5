"""Child process of test_gpu_boundary.py::test_ipc_gather_from_a_second_process: opens the parent's frame planes through CUDA IPC
(atray_ipc_open) and renders the bands of `rank` of `world` straight into them, as a peer rank of bench.py does.
usage: python tests/ipc_child.py mesh.obj rescale W H rank world <hex handle bgrx> <hex handle t> <hex handle prim>"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import atray_b200 as A  # noqa: E402


def main():
    path, rescale, W, H, rank, world = sys.argv[1], float(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6])
    handles = [bytes.fromhex(h) for h in sys.argv[7:10]]
    m = A.load_model_data(path, 4)
    A.place_default(m, rescale)
    sc = A.Scene(m)
    A.prep_scene(sc)
    ctx = sc.ctx
    planes = [ctx.ipc_open(h) for h in handles]
    ctx.render_device(A.set_camera(W, H), ctx.opts(rank=rank, world=world, band_rows=8), *planes)
    ctx.sync()
    for p in planes:
        ctx.ipc_close(p)
    print("child rank %d of %d done" % (rank, world))


if __name__ == "__main__":
    main()
