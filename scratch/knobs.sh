#!/bin/bash
# usage: knobs.sh "ENV1=.. ENV2=.." ...   -> sweep2 time at ne120 for each setting
for cfg in "$@"; do
  echo "== $cfg"
  env $cfg timeout 120 python profiles/prof_run.py 120 72 40 2>&1 | grep "sweep2" | tr '\n' ' '; echo
done
