cd /root/repo
for p in 1 2; do for bn in 128 256; do
echo "== pair=$p bn=$bn"
ADAPT_PAIR=$p KC_BN=$bn KC_ONLY=conv4_1x1 timeout 300 python tools/kernel_check.py --group trace 2>&1 | grep grid | cut -c1-400
done; done
