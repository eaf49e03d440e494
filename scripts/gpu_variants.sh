#!/bin/bash
# usage: gpu_variants.sh TAG "libtag1 libtag2 ..." "N1 N2 ..." [extra bench args]
TAG=$1; LIBS=$2; NS=$3; shift 3
mkdir -p gpurun_out
for lib in $LIBS; do
  if [ "$lib" = "default" ]; then unset W3D_CUDA_LIB; else export W3D_CUDA_LIB=mpi_cuda_b200/_native/libwave3d_b200.$lib.so; fi
  for n in $NS; do
    bc=ddp; [ "$n" = "512" ] && bc=ddd
    out=gpurun_out/${TAG}_${lib}_$n
    timeout 200 python bench.py --N $n --bc $bc --no-e2e "$@" > $out.json 2> $out.err
    python - <<PY
import json
try:
    d = json.load(open("$out.json"))
    x = d["exchange"]
    print("$lib $n %.5f ms/step %.1f GLUP/s valid=%s wait %.2f copy %.2f us/pass | %s" % (d["ms_per_step"], d["value"], d.get("validated_vs_oracle"), 2e3*x["flag_wait_ms_per_step"], 2e3*x["halo_copy_ms_per_step"], d["plan"].split("||")[1][:40]))
except Exception as e:
    print("$lib $n failed", e, open("$out.err").read()[-400:])
PY
  done
done
