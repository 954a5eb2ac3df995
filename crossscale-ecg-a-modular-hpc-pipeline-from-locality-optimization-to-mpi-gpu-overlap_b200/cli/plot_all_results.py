"""`python -m crossscale.cli.plot_all_results [--results-dir D --csv F --batch B]` — see ``cli/plots.py``.
Reference: ``Module_1/plot_all_results.py:1-129`` (merged A0-A4 CSV, amortised "effective" A4 throughput ``:56-58``)."""
import sys

from .plots import main as _main


def main(argv=None):
    return _main(["all"] + list(sys.argv[1:] if argv is None else argv))


if __name__ == "__main__":
    main()
