"""`python -m crossscale.cli.plot_part2 [--results-dir D --csv F --batch B]` — see ``cli/plots.py``.
Reference: ``Module_2/plot_part2.py:1-17`` and the two figures of ``Module_2/benchmark_part_2.py:141-178``."""
import sys

from .plots import main as _main


def main(argv=None):
    return _main(["part2"] + list(sys.argv[1:] if argv is None else argv))


if __name__ == "__main__":
    main()
