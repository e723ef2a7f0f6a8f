//go:build !icicle

package icicle_bn254

import (
	"fmt"

	"github.com/consensys/gnark/backend"
	groth16_bn254 "github.com/consensys/gnark/backend/groth16/bn254"
	"github.com/consensys/gnark/backend/witness"
	cs "github.com/consensys/gnark/constraint/bn254"
)

// HasIcicle is false when the package is built without the icicle tag (groth16.go:200-204 dispatches on it).
const HasIcicle = false

// Prove reports that the accelerated prover was not compiled in.
func Prove(r1cs *cs.R1CS, pk *ProvingKey, fullWitness witness.Witness, opts ...backend.ProverOption) (*groth16_bn254.Proof, error) {
	return nil, fmt.Errorf("icicle backend requested but program compiled without 'icicle' build tag")
}
