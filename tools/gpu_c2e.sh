#!/bin/bash
for seg in 32 64 128 256 512; do echo "seg=$seg"; SPMMDB_CSR_SEG=$seg timeout 300 python tools/time_kernels.py --cases c2,c2l 2>/dev/null | cut -c1-170; done
