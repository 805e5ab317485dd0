#!/bin/bash
for lc in 2 4 8; do echo "long_ctas=$lc"; SPMMDB_CSR_LONG_CTAS=$lc timeout 300 python tools/time_kernels.py --cases c2,c2l 2>/dev/null | cut -c1-170; done
echo "seg=32 lc=4"; SPMMDB_CSR_SEG=32 SPMMDB_CSR_LONG_CTAS=4 timeout 300 python tools/time_kernels.py --cases c2,c2l 2>/dev/null | cut -c1-170
echo "seg=128 lc=4"; SPMMDB_CSR_SEG=128 SPMMDB_CSR_LONG_CTAS=4 timeout 300 python tools/time_kernels.py --cases c2,c2l 2>/dev/null | cut -c1-170
