O=gpurun_out/r2exp14; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
for upb in 5 8 10 12 15 16; do
  echo "barrier kernel, UPB=$upb"
  BMHRL_RNN_EXCHANGE=0 BMHRL_RNN_UPB=$upb timeout 300 python tools/trace_step.py > $O/trace_upb$upb.txt 2> $O/trace_upb$upb.err; head -1 $O/trace_upb$upb.txt; grep rnn_layer $O/trace_upb$upb.txt | head -2
done
echo "no critic at all"
BMHRL_DIAG_NO_CRITIC=1 timeout 300 python tools/trace_step.py > $O/trace_nocritic.txt 2> $O/trace_nocritic.err; head -1 $O/trace_nocritic.txt
