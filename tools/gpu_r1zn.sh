#!/bin/bash
cd /root/repo; T=r1zn; O=gpurun_out; mkdir -p $O
timeout 300 python tools/bench_extras.py > $O/${T}_extras.jsonl 2> $O/${T}_extras.err; echo "extras rc=$?"; tail -3 $O/${T}_extras.jsonl; tail -3 $O/${T}_extras.err
