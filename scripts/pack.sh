#!/bin/sh
# Source tarball of the tracked tree (reference: /root/reference/scripts/pack.sh).
set -e
cd "$(dirname "$0")/.."
out=${1:-kungfu-b200-$(git rev-parse --short HEAD).tar.gz}
git archive --format=tar.gz --prefix=kungfu-b200/ -o "$out" HEAD
echo "$out"
