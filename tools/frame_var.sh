# run-to-run spread of the whole-frame figure, knobs alternating
for i in 1 2 3 4; do
  for cfg in ${CFGS:-"X=1" "PL_NO_PDL=1" "PL_BLUR_FORM=reg"}; do
    echo -n "$cfg  "; env $cfg python tools/frame_exp.py child
  done
done
