for nb in 384 512 640 768; do for mdl in 1.0 1.25 1.5; do
echo "NB=$nb MODEL=$mdl: $(LA_B200_LU_NB=$nb LA_B200_LU_MODEL=$mdl python tools/factor_once.py getrf 16384 3 | tail -1)"; done; done
echo "NO_LOOKAHEAD: $(LA_B200_NO_LOOKAHEAD=1 python tools/factor_once.py getrf 16384 3 | tail -1)"
for nbo in 512 768 1024 1536; do for s in 4 8 16; do
echo "CHOL NBO=$nbo SIDE=$s: $(LA_B200_CHOL_NBO=$nbo LA_B200_CHOL_SIDE_SMS=$s python tools/factor_once.py potrf 16384 3 | tail -1)"; done; done
