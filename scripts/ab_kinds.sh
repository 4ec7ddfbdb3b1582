#!/bin/bash
for so in "$@"; do echo "== $so"; KOROLLUP_LIB=$PWD/$so timeout 200 python scripts/diag_kinds.py 2>&1 | tail -7; done
