for v in tools/_var_qy8.so "" tools/_var_qy16.so; do
  if [ -n "$v" ]; then export M3R_B200_LIB=$PWD/$v; else unset M3R_B200_LIB; fi
  timeout 300 python tools/refine_var.py 2>&1 | tail -1
done
