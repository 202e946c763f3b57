#!/bin/bash
# Run under gpurun (one B200): decode-step probe of the tuning builds of the int4 GEMV
# (build them first: python lit-llama-ja_b200/build.py --variant <name> <defines>).
#   profiles/variant_sweep.sh "name:lib:cfg:ENV=VAL ..."   lib = '' for the default library, cfg = '' for the default WxB,
#   the optional 4th field is one more environment assignment for that run
mkdir -p gpurun_out
L=$PWD/lit-llama-ja_b200/lit_llama
LOG=gpurun_out/sweep.log
: > $LOG
export PROBE_SEQS=${PROBE_SEQS:-q,p,f,d,qpfd}
for spec in ${1:-base::}; do
  IFS=: read -r name lib cfg extra <<< "$spec"
  echo "== $name (lib=${lib:-default} cfg=${cfg:-default})" >> $LOG
  env LLB200_LIB=${lib:+$L/$lib} LLB200_GEMV_CFG=$cfg ${extra//,/ } timeout -s KILL 200 python profiles/step_probe.py > gpurun_out/sweep_$name.out 2>&1
  echo "rc=$?" >> $LOG
  grep "cfg=" gpurun_out/sweep_$name.out >> $LOG || tail -5 gpurun_out/sweep_$name.out >> $LOG
done
cat $LOG
