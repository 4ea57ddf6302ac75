for t in 4 8 16 32; do python tools/prof_run.py pigs 75776 $t 2 1 | tail -1; done
for t in 2 4 8 16; do python tools/prof_run.py hepar2 265216 $t 2 1 | tail -1; done
