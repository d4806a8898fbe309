cd distributed-neural-networks_b200 && python build.py >/dev/null 2>&1; cd ..
for v in "B2P_FC_FUSED=1" "B2P_FC_FUSED=1 B2P_FC_ONLY=1" "B2P_FC_FUSED=1 B2P_FC_ONLY=2" "B2P_FC_FUSED=0" "B2P_FC_FUSED=1 B2P_NO_PDL=1" "B2P_FC_FUSED=0 B2P_NO_PDL=1"; do
  echo "--- $v"; env $v timeout 300 python scripts/time_stages_graph.py tf32x3 1 128 1024 2>&1 | grep part1
done
