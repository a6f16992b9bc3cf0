set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
( time python -m fedb200 custom --clients 3 --rounds 300 --quiet ) 2>&1 | tail -6
( time python -m fedb200 custom --clients 8 --rounds 300 --quiet --precision fp32 ) 2>&1 | tail -6
( time python -m fedb200 sklearn-avg --clients 4 --rounds 5 --quiet ) 2>&1 | tail -12
( time python -m fedb200 sklearn-avg --clients 4 --rounds 3 --quiet --policy reference_faithful ) 2>&1 | tail -10
( time python -m fedb200 tune --clients 4 --quiet ) 2>&1 | tail -22
