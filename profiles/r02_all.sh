#!/bin/bash
# every piece of round-2 evidence in one GPU call (all ncu passes run after their plain command exited 0)
bash profiles/r02_capture.sh
bash profiles/r02_ncu_s1.sh
bash profiles/r02_launches.sh
