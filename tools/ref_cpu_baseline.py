#!/usr/bin/env python
"""CPU baseline of the junction-tree path with the UNMODIFIED reference (oracle/_ref), as a separate tool: what the
reference's `BayesianNetwork -a 2 -t N -f0 .. -f3 .. -f4 ..` spends in its `jt` timer, for -t 1 and -t N.

  python tools/ref_cpu_baseline.py <bn.xml> <test cases (LIBSVM)> <potential tables> [threads ...]

Never linked into the product: the CLI (fastbn_b200/bin/BayesianNetwork) accepts -t like the reference and ignores it."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refdrv  # noqa: E402


def main():
    if len(sys.argv) < 4:
        raise SystemExit(__doc__)
    xml, cases, pt = sys.argv[1:4]
    threads = [int(x) for x in sys.argv[4:]] or [1, os.cpu_count() or 1]
    if not refdrv.available():
        raise SystemExit("oracle/_ref/libfastbn_ref.so is missing: run oracle/build_ref.sh where /root/reference is present")
    r = refdrv.Reference(xml, cases)
    if r.n_cases < 20:
        raise SystemExit("the reference's driver loop needs at least 20 test cases (progress % (n / 20))")
    for t in threads:
        acc, secs = r.predict_timed(pt, t)
        print("-t %d: jt %.4f s, %.1f cases/s, accuracy = %g" % (t, secs, r.n_cases / secs, acc))
    r.close()


if __name__ == "__main__":
    main()
