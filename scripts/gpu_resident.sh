#!/bin/bash
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02h_pytest.log 2>&1; tail -4 gpurun_out/r02h_pytest.log
timeout 120 titan_b200/bin/titan_facade_dump bench c1 10000 2>&1 | grep '^{'
timeout 120 titan_b200/bin/titan_facade_dump bench lattice10 10000 2>&1 | grep '^{'
TB_RESIDENT=0 timeout 120 titan_b200/bin/titan_facade_dump bench c1 10000 2>&1 | grep '^{'
TB_RESIDENT=0 timeout 120 titan_b200/bin/titan_facade_dump bench lattice10 10000 2>&1 | grep '^{'
