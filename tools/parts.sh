#!/bin/bash
for P in 6 8 10; do echo "antsafe parts=$P: $(CMBPO_POST_PARTS_RT=$P python tools/fwd_bench.py 1000000 10 antsafe_step_1M 2>&1 | tail -1 | sed 's/.*whole step/whole step/')"; done
for P in 6 9 ; do echo "halfcheetah parts=$P: $(CMBPO_POST_PARTS_RT=$P python tools/fwd_bench.py 1000000 10 halfcheetah_schedule_L10_1M 2>&1 | tail -1 | sed 's/.*whole step/whole step/')"; done
for P in 6 8 10; do echo "antsafe 100k parts=$P: $(CMBPO_POST_PARTS_RT=$P python tools/fwd_bench.py 100000 20 antsafe_step_1M 2>&1 | tail -1 | sed 's/.*whole step/whole step/')"; done
