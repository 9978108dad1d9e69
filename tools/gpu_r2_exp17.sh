O=gpurun_out/r2exp17; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
BMHRL_TIMELINE=1 timeout 300 python tools/trace_step.py > $O/trace.txt 2> $O/trace.err; head -3 $O/trace.txt
BMHRL_DIAG_NO_CRITIC=1 timeout 300 python tools/trace_step.py > $O/trace_nocritic.txt 2> $O/trace_nocritic.err; head -1 $O/trace_nocritic.txt
timeout 300 python tools/microbench16.py > $O/microbench16.jsonl 2> $O/mb.err; cat $O/microbench16.jsonl | cut -c1-200
