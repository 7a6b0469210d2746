"""CPU: the algorithmic bytes / flops bench.py divides by the measured kernel times (`roofline.achieved`, the `kernels`
tables) against SURVEY section 8(d)'s per-unit figures, with the argument tuples built from the header prototypes by
parameter NAME — a kernel_work() that reads the wrong positional argument after an ABI change fails here."""
import importlib.util
from pathlib import Path

from tests.test_abi import _prototypes

ROOT = Path(__file__).resolve().parent.parent
spec = importlib.util.spec_from_file_location("nkbs_bench", ROOT / "bench.py")
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)
PROTOS = _prototypes()


def _args(name, **named):
    """Positional argument tuple of entry point `name`: pointers are a truthy placeholder unless named None."""
    out = []
    for decl in PROTOS[name]:
        pname = decl.replace("*", " ").split()[-1]
        if pname in named:
            out.append(named[pname])
        elif "*" in decl:
            out.append(1)  # some non-null pointer
        else:
            out.append(0)
    assert not set(named) - {d.replace("*", " ").split()[-1] for d in PROTOS[name]}, "unknown parameter name"
    return tuple(out)


def test_encode_bytes_per_sample_match_survey():
    # main grid L=16, F=2: 1164 B/sample; proposal grid L=5: 372 B/sample
    for L, per in ((16, 1164), (5, 372)):
        kind, amount = bench.kernel_work("nkbs_encode_fwd", _args("nkbs_encode_fwd", R=4096, S=48, L=L, F=2))
        assert kind == "hbm" and amount == 4096 * 48 * per
        kind, amount = bench.kernel_work("nkbs_encode_bwd", _args("nkbs_encode_bwd", N=1000, L=L, F=2))
        assert amount == 1000 * (12 + 4 * L * 2 + 8 * L * 2 * 4)


def test_composite_bytes_per_ray_match_survey():
    # S=48, C=19, every input and output present: 5092 B/ray
    a = _args("nkbs_composite_fwd", R=100, S=48, C=19, weights_in=None)
    kind, amount = bench.kernel_work("nkbs_composite_fwd", a)
    assert kind == "hbm" and amount == 100 * 5092


def test_adam_and_mlp_work():
    numel = [1000, 24]
    a = _args("nkbs_adam_step", num_tensors=2, numel=numel, zero_grads=1)
    assert bench.kernel_work("nkbs_adam_step", a) == ("hbm", 32 * 1024)
    a = _args("nkbs_adam_step", num_tensors=2, numel=numel, zero_grads=0)
    assert bench.kernel_work("nkbs_adam_step", a) == ("hbm", 28 * 1024)
    # base MLP 32 -> 64 -> 16: 3072 MAC/sample; backward = data + weight gradient contractions
    a = _args("nkbs_mlp_tc_fwd", N=10, nl=2, dims=[32, 64, 16])
    assert bench.kernel_work("nkbs_mlp_tc_fwd", a) == ("tensor", 10 * 3072 * 2)
    a = _args("nkbs_mlp_tc_bwd", N=10, nl=2, dims=[32, 64, 16])
    assert bench.kernel_work("nkbs_mlp_tc_bwd", a) == ("tensor", 10 * 3072 * 4)
    # colour head 63 -> 64 -> 64 -> 3: 8320 MAC/sample
    a = _args("nkbs_head_tc_fwd", N=10, nl=3, dims=[63, 64, 64, 3])
    assert bench.kernel_work("nkbs_head_tc_fwd", a) == ("tensor", 10 * 8320 * 2)


def test_density_and_scatter_work_follow_the_named_arguments():
    a = _args("nkbs_density_fwd", R=7, S=96, L=5)
    assert bench.kernel_work("nkbs_density_fwd", a) == ("hbm", 7 * 96 * (8 + 8 * 5 * 2 * 4 + 4))
    a = _args("nkbs_density_fwd_save", R=7, S=96, L=5)
    assert bench.kernel_work("nkbs_density_fwd_save", a) == ("hbm", 7 * 96 * (8 + 320 + 4 + 40 + 16))
    a = _args("nkbs_density_bwd_saved", N=1234, L=5)
    assert bench.kernel_work("nkbs_density_bwd_saved", a) == ("hbm", 1234 * 88)
    a = _args("nkbs_encode_bwd_rays", R=7, S=96, L=5)
    assert bench.kernel_work("nkbs_encode_bwd_rays", a) == ("hbm", 7 * 96 * (8 + 40 + 320))
    a = _args("nkbs_fixed_point_flush", n=1000)
    assert bench.kernel_work("nkbs_fixed_point_flush", a) == ("hbm", 8000)
    # every entry point the step launches has a roofline entry (no silent zero)
    for name in ("nkbs_initial_sampler", "nkbs_weights_resample_fwd", "nkbs_ray_losses_weighted", "nkbs_composite_bwd",
                 "nkbs_head_tc_bwd"):
        assert name in PROTOS
