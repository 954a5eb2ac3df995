"""`python -m crossscale.cli.plot_locality [--results-dir D --csv F --batch B]` — see ``cli/plots.py``.
Reference: ``Module_1/plot_locality.py:7-73`` (throughput vs batch, stacked data/h2d/compute bars at ``--batch``)."""
import sys

from .plots import main as _main


def main(argv=None):
    return _main(["locality"] + list(sys.argv[1:] if argv is None else argv))


if __name__ == "__main__":
    main()
