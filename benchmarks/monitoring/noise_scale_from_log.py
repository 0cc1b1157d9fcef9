#!/usr/bin/env python
"""Scrape gradient-noise-scale readings out of training logs into CSV and a terminal plot.

    python benchmarks/monitoring/noise_scale_from_log.py run.log [more.log ...] [--csv out.csv]

Recognises the lines examples/elastic_resnet_gns.py and the GNS optimizer print
("step 120 ... noise scale 1834.2 ..."). Reference: /root/reference/experimental/gradient_noise/
(log-scraping plot of the noise scale)."""
import argparse
import re
import sys

PAT = re.compile(r'step\s+(\d+).*?noise[ _-]scale[^0-9+-]*([-+]?[0-9]*\.?[0-9]+(?:[eE][-+]?\d+)?)', re.I)


def scrape(lines):
    out = []
    for line in lines:
        m = PAT.search(line)
        if m:
            out.append((int(m.group(1)), float(m.group(2))))
    return out


def ascii_plot(points, width=60, height=12):
    if not points:
        return '(no readings)'
    ys = [y for _, y in points]
    lo, hi = min(ys), max(ys)
    span = (hi - lo) or 1.0
    rows = [[' '] * width for _ in range(height)]
    for i, (_, y) in enumerate(points):
        c = min(width - 1, i * width // len(points))
        r = height - 1 - int((y - lo) / span * (height - 1))
        rows[r][c] = '*'
    return '\n'.join('%10.3g |%s' % (hi - k * span / (height - 1), ''.join(r)) for k, r in enumerate(rows))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('logs', nargs='+')
    ap.add_argument('--csv', default='')
    args = ap.parse_args()
    pts = []
    for p in args.logs:
        with open(p, errors='replace') as f:
            pts += scrape(f)
    pts.sort()
    if args.csv:
        with open(args.csv, 'w') as f:
            f.write('step,noise_scale\n')
            f.writelines('%d,%g\n' % p for p in pts)
    print(ascii_plot(pts))
    print('%d readings' % len(pts))


if __name__ == '__main__':
    sys.exit(main())
