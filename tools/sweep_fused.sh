for pdl in 0 1; do for mode in 3 0; do
  echo -n "pdl=$pdl mode=$mode: "
  GGML_B200_FUSED_PDL=$pdl GGML_B200_FUSED_DBG_MODE=$mode python tools/mmvq_probe.py q4_K 32000 4096 9 10 | grep -o "copies.*"
done; done
