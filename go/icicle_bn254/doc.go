// Package icicle_bn254 is the cgo host shim that puts libg16b200 (a B200-native Groth16 proving core) behind
// gnark's accelerated-prover surface: groth16.Prove(r1cs, pk, fullWitness, backend.WithIcicleAcceleration()).
//
// It replaces package backend/groth16/bn254/icicle of ingonyama-zk/gnark-icicle (same package name, same
// exported identifiers: HasIcicle, ProvingKey, Setup, DummySetup, Prove) and binds include/g16b200.h instead of
// github.com/ingonyama-zk/iciclegnark.  NOTE: neither build machine of this repository has a Go toolchain, so
// this package is provided as reviewed-but-uncompiled source; every line of logic that matters lives behind the
// C ABI, where it is tested (tests/ drives the same ABI through ctypes).
package icicle_bn254
