"""paritytech-bn_b200: B200-native batched BN254 pairing engine (import as `paritytech_bn_b200`)."""
