for n in 8192 16384 32768; do for nbo in 512 768 1024; do for s in 4 8; do
echo "N=$n NBO=$nbo SIDE=$s: $(LA_B200_CHOL_NBO=$nbo LA_B200_CHOL_SIDE_SMS=$s python tools/factor_once.py potrf $n 3 | tail -1)"; done; done; done
