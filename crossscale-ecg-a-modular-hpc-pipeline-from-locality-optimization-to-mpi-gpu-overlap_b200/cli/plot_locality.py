"""`python -m crossscale.cli.plot_locality [--results-dir D --csv F --batch B]` — see ``cli/plots.py``."""
import sys

from .plots import main as _main


def main(argv=None):
    return _main(["locality"] + list(sys.argv[1:] if argv is None else argv))


if __name__ == "__main__":
    main()
