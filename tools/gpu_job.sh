for v in "" tools/_var_acc_192_3.so tools/_var_acc_288_2.so tools/_var_acc_128_4.so tools/_var_acc_160_3.so tools/_var_acc_128_5.so tools/_var_acc_160_4.so tools/_var_acc_96_6.so; do
  if [ -n "$v" ]; then export M3R_B200_LIB=$PWD/$v; else unset M3R_B200_LIB; fi
  KINDS=rays,calib timeout 300 python tools/gn_var.py 2>&1 | tail -2
done
