package icicle_bn254

import (
	"sync"

	groth16_bn254 "github.com/consensys/gnark/backend/groth16/bn254"
	cs "github.com/consensys/gnark/constraint/bn254"
)

// deviceInfo is what stays resident on the GPU between proofs: a libg16b200 context and the handle of the
// uploaded key (bases in Montgomery affine form, index maps, domain twiddles).  It is never serialised, so a key
// read back with ReadFrom re-uploads lazily on its next Prove.
type deviceInfo struct {
	mu         sync.Mutex // the library context is thread-compatible (one call in flight): Prove and the bsb22 hint hold it
	ctx        ctxHandle
	handle     uint64
	commitKeys []uint64 // one resident pedersen.ProvingKey per BSB22 commitment
}

// ProvingKey embeds the native key so that every (un)marshalling method of groth16.ProvingKey is inherited.
type ProvingKey struct {
	groth16_bn254.ProvingKey
	*deviceInfo
}

// Setup and DummySetup run on the CPU exactly as in gnark; only proving is accelerated.
func Setup(r1cs *cs.R1CS, pk *ProvingKey, vk *groth16_bn254.VerifyingKey) error {
	return groth16_bn254.Setup(r1cs, &pk.ProvingKey, vk)
}

func DummySetup(r1cs *cs.R1CS, pk *ProvingKey) error {
	return groth16_bn254.DummySetup(r1cs, &pk.ProvingKey)
}
