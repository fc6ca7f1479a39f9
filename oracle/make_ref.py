"""Recipe for `baseline/_ref/`: the files of the UNMODIFIED reference that the hot path consists of,
placed where they can travel to the GPU box.

TEST / BENCH INFRASTRUCTURE.  `/root/reference` does not exist on the GPU box and reference sources are
never committed; `baseline/_ref/` is git-ignored but NOT gpurun-ignored (like the built `libiqlb.so`),
so a copy made here ships with the snapshot.  It is used for exactly three things:

  * `bench.py --impl reference`: the reference's own `ReplayBuffer.sample()` + `ImplicitQLearning.train()`
    (algorithms/offline/iql.py:211-221, 639-662) timed on the box's host cores;
  * `bench.py`'s `reference_cuda` key: the same stock-PyTorch code on the same B200 (1 agent, eager);
  * `-m gpu` parity tests of the bf16 step against the reference run under CUDA autocast (O-ac-cuda,
    SURVEY.md 8c), and `scripts/gen_cuda_golden.py`, which writes the committed CUDA fixtures.

Everything degrades to the committed fixtures / the oracle port when `baseline/_ref/` is absent.

    python -m oracle.make_ref            # copy (idempotent); called by __graft_entry__.build()
"""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.environ.get("IQLPREF_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(ROOT, "baseline", "_ref")

# (path under the reference root, is_directory)
FILES = (
    ("algorithms/offline/iql.py", False),        # the hot path (as shipped: bf16 autocast)
    ("algorithms/minari/iql.py", False),         # the reference's own fp32 statement of the same step
    ("gp_reward-priors/optbnn", True),           # RewardMLP / RewardPT / trainers imported by iql.py:27-33
)


def make_ref(verbose: bool = True) -> bool:
    """Copy the hot-path files of the reference into baseline/_ref/.  Returns False (and does nothing)
    when the reference is not present (e.g. on the GPU box, where the shipped copy is used as is)."""
    if not os.path.isfile(os.path.join(SRC, FILES[0][0])):
        if verbose:
            print(f"make_ref: no reference under {SRC}; leaving {DST} as it is", file=sys.stderr)
        return False
    for rel, is_dir in FILES:
        s, d = os.path.join(SRC, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        if is_dir:
            if os.path.isdir(d):
                shutil.rmtree(d)
            shutil.copytree(s, d, ignore=shutil.ignore_patterns("__pycache__", "*.pyc", "*.ipynb", "*.h5", "*.hdf5",
                                                                "*.pt", "*.png", "*.pdf"))
        else:
            shutil.copyfile(s, d)
    with open(os.path.join(DST, "README"), "w") as f:
        f.write("Unmodified files of ml4ai/iqlpref copied by oracle/make_ref.py (git-ignored; not product source).\n")
    if verbose:
        print(f"make_ref: {DST} populated from {SRC}", file=sys.stderr)
    return True


if __name__ == "__main__":
    make_ref()
