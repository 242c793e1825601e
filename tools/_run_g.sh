timeout 300 python -m pytest tests/test_ops_gpu.py -m gpu -q -p no:cacheprovider -k "benchmark_variants or batchnorm_keeps or instnorm or unet_matches" 2>&1 | tail -8
for l in unet_256_256_48 unet_512_512_24 unet_1024_1024_12 unet_2048_2048_6 unet_256_512_24; do
  echo "1cta: $(TGB_FPROP_2CTA=0 timeout 60 build/selftest $l | grep -v wgrad | grep -v selftest | tr '\n' ' ')"
  echo "2cta: $(TGB_FPROP_2CTA=1 timeout 60 build/selftest $l | grep -v wgrad | grep -v selftest | tr '\n' ' ')"
done
