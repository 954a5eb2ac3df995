"""`python -m crossscale.cli.plot_part3 [--results-dir D --csv F --batch B]` — see ``cli/plots.py``.
Reference: ``Module_3/plot_part3.py:1-149`` (pseudo-FL) and ``Module_3/TRUE_FL_M3/plot_part3.py:1-164`` (FedAvg; its glob
``fedavg_results_*.csv`` does not match the file the driver writes — here both names are read)."""
import sys

from .plots import main as _main


def main(argv=None):
    return _main(["part3"] + list(sys.argv[1:] if argv is None else argv))


if __name__ == "__main__":
    main()
