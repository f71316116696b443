# developer probe: each case in its own process under `timeout`, so that a hung kernel costs seconds
run() { echo "== $*"; timeout 40 env B200_AWQ_DEBUG=1 "$@" > /tmp/p.out 2>&1; echo "rc=$? $(tail -1 /tmp/p.out)"; }
run env B200_AWQ_SPLITS=1 B200_AWQ_NT=1 python tools/awq_probe.py 128 256 2048
run env B200_AWQ_SPLITS=2 B200_AWQ_NT=1 python tools/awq_probe.py 128 256 2048
run env B200_AWQ_SPLITS=1 B200_AWQ_NT=2 python tools/awq_probe.py 128 512 2048
run env B200_AWQ_SPLITS=1 B200_AWQ_NT=3 python tools/awq_probe.py 128 768 2048
run env B200_AWQ_SPLITS=2 B200_AWQ_NT=3 python tools/awq_probe.py 128 768 2048
run env B200_AWQ_SPLITS=3 B200_AWQ_NT=2 python tools/awq_probe.py 200 1024 3072
run python tools/awq_probe.py 128 3584 5120
run python tools/awq_probe.py 128 27648 5120
run python tools/awq_probe.py 1 5120 13824
run python tools/awq_probe.py 64 5120 2560
