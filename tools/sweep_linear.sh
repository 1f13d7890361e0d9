# span (first to last in-kernel timestamp) of one fused Linear launch under different shared-memory plans
for shape in "360 180 resid" "180 180 aggr" "192 180 resid"; do
  python tools/trace_gemm.py $shape 2>&1 | tail -1
  for ops in 2 3 4; do for bufs in 1 2; do for res in 0 1; do
    HATB200_G3_OPS=$ops HATB200_G3_STGBUFS=$bufs HATB200_G3_RESIDENT=$res python tools/trace_gemm.py $shape 2>&1 | tail -1
  done; done; done
done
python tools/trace_gemm.py 180 720 plain; python tools/trace_gemm.py 180 540 plain; python tools/trace_gemm.py 180 360 plain
