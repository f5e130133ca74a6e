set -x
timeout 300 python tools/team_probe.py --n 32768 --launches 50 --ks 8 --checks 6 2>&1 | grep -v Warning
timeout 300 python tools/team_probe.py --n 32768 --launches 50 --ks 7,6,5 --ariths strict-fma --clusters 1 --checks 1 --tag _k765 2>&1 | grep -v Warning
timeout 300 python tools/team_probe.py --n 16384 --launches 100 --ks 8 --ariths strict-fma,fp32 --clusters 1 --checks 1 2>&1 | grep -v Warning
timeout 300 python tools/team_probe.py --n 8192 --launches 200 --ks 8 --ariths strict-fma --clusters 1 --checks 1 2>&1 | grep -v Warning
