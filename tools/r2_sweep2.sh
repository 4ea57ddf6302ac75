cd /root/repo
for cb in 2 4; do for t in 4 8; do echo "== hepar2 team $t cb $cb"; FBN_CB=$cb python tools/prof_run.py hepar2 200000 $t 2 1 | tail -1; done; done
echo "== hepar2 level"; python tools/prof_run.py hepar2 200000 0 2 2 | tail -1
for cb in 2 4; do for t in 8 16; do echo "== pigs team $t cb $cb"; FBN_CB=$cb python tools/prof_run.py pigs 60000 $t 2 1 | tail -1; done; done
echo "== pigs level"; python tools/prof_run.py pigs 60000 0 2 2 | tail -1
echo "== pigs level res 128"; FBN_LEVEL_RESIDENT=128 python tools/prof_run.py pigs 60000 0 2 2 | tail -1
echo "== munin_small level"; python tools/prof_run.py munin_small 8192 0 2 2 | tail -1
echo "== munin fused team 16"; python tools/prof_run.py munin 4096 16 2 1 | tail -1
