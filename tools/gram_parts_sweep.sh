big="1000000 500 30"
for sbp in 128 256 512 1024; do DVF_GRAM_SBP=$sbp DVF_GRAM_DEBUG=1 python tools/gram_tune.py $big 2>&1 | tail -3; done
DVF_GRAM_LAG=1 python tools/gram_tune.py $big | tail -1
DVF_GRAM_LAG=1 DVF_GRAM_SBP=256 python tools/gram_tune.py $big | tail -1
DVF_GRAM_PARTS=0 python tools/gram_tune.py $big | tail -1
DVF_GRAM_DEBUG=1 python tools/gram_tune.py 1000000 1000 50 2>&1 | tail -3
DVF_GRAM_DEBUG=1 DVF_GRAM_PARTS=1 python tools/gram_tune.py 1000000 1000 50 2>&1 | tail -3
DVF_GRAM_DEBUG=1 python tools/gram_tune.py 1000000 100 30 2>&1 | tail -3
DVF_GRAM_DEBUG=1 python tools/gram_tune.py 100000 100 30 2>&1 | tail -3
DVF_GRAM_DEBUG=1 python tools/gram_tune.py 999983 500 30 2>&1 | tail -3
DVF_GRAM_DEBUG=1 python tools/gram_tune.py 1000000 200 50 2>&1 | tail -3
DVF_GRAM_DEBUG=1 python tools/gram_tune.py 20000 500 30 2>&1 | tail -3
