#!/bin/bash
for so in "$@"; do
  echo "$so: $(KOROLLUP_LIB=$PWD/$so timeout 200 python scripts/diag_one.py 0.5 0.95 0.99 2>&1 | tail -1)"
done
