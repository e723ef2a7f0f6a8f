//go:build icicle

package icicle_bn254

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -L${SRCDIR}/../../gnark_icicle_b200 -lg16b200 -Wl,-rpath,${SRCDIR}/../../gnark_icicle_b200
#include <stdlib.h>
#include "g16b200.h"
*/
import "C"

import (
	"fmt"
	"math/big"
	"os"
	"runtime"
	"strconv"
	"time"
	"unsafe"

	curve "github.com/consensys/gnark-crypto/ecc/bn254"
	"github.com/consensys/gnark-crypto/ecc/bn254/fr"
	"github.com/consensys/gnark/backend"
	groth16_bn254 "github.com/consensys/gnark/backend/groth16/bn254"
	"github.com/consensys/gnark/backend/witness"
	"github.com/consensys/gnark-crypto/ecc"
	"github.com/consensys/gnark-crypto/ecc/bn254/fr/hash_to_field"
	"github.com/consensys/gnark/constraint"
	cs "github.com/consensys/gnark/constraint/bn254"
	"github.com/consensys/gnark/constraint/solver"
	fcs "github.com/consensys/gnark/frontend/cs"
	"github.com/consensys/gnark/logger"
)

// HasIcicle tells the groth16 facade that the accelerated prover is compiled in.
const HasIcicle = true

type ctxHandle = *C.g16_ctx

func lastError(ctx ctxHandle, what string) error {
	return fmt.Errorf("%s: %s", what, C.GoString(C.g16_last_error(ctx)))
}

func boolBytes(b []bool) []byte {
	out := make([]byte, len(b))
	for i, v := range b {
		if v {
			out[i] = 1
		}
	}
	return out
}

func ptrOf[T any](s []T) unsafe.Pointer {
	if len(s) == 0 {
		return nil
	}
	return unsafe.Pointer(&s[0])
}

// setupDevicePointers uploads the key once.  gnark-crypto elements are already the memory image the library
// expects (little-endian Montgomery limbs), so slices are handed over as they are -- no conversion, no
// stripping of infinity points.
//
// commitmentInfo decides which wires reach the K MSM (prove.go:241-246): W[nbPublic:] without the privately committed
// wires and the commitment wires.  The reference filters the wire vector on the host for every proof (filterHeap);
// here the surviving wire indices become part of the resident key (k_index) and the device gathers.
func (pk *ProvingKey) setupDevicePointers(nbPublic int, commitmentInfo constraint.Groth16Commitments) error {
	if pk.deviceInfo != nil {
		return nil
	}
	var kIndex []uint32
	if len(commitmentInfo) != 0 {
		nbWires := len(pk.InfinityA)
		removed := make([]bool, nbWires)
		for _, c := range commitmentInfo {
			for _, w := range c.PrivateCommitted {
				removed[w] = true
			}
			removed[c.CommitmentIndex] = true
		}
		kIndex = make([]uint32, 0, len(pk.G1.K))
		for w := nbPublic; w < nbWires; w++ {
			if !removed[w] {
				kIndex = append(kIndex, uint32(w))
			}
		}
		if len(kIndex) != len(pk.G1.K) {
			return fmt.Errorf("commitment info leaves %d private wires, pk.G1.K has %d", len(kIndex), len(pk.G1.K))
		}
	}
	ctx, err := newContext()
	if err != nil {
		return err
	}
	infA, infB := boolBytes(pk.InfinityA), boolBytes(pk.InfinityB)
	var pin runtime.Pinner
	defer pin.Unpin()
	for _, p := range []unsafe.Pointer{ptrOf(pk.G1.A), ptrOf(pk.G1.B), ptrOf(pk.G1.K), ptrOf(pk.G1.Z), ptrOf(pk.G2.B), ptrOf(infA), ptrOf(infB), ptrOf(kIndex)} {
		if p != nil {
			pin.Pin(p)
		}
	}
	var d C.g16_pk_desc
	d.domain_n = C.uint64_t(pk.Domain.Cardinality)
	d.g1_a, d.n_a = ptrOf(pk.G1.A), C.uint64_t(len(pk.G1.A))
	d.g1_b, d.n_b = ptrOf(pk.G1.B), C.uint64_t(len(pk.G1.B))
	d.g1_k, d.n_k = ptrOf(pk.G1.K), C.uint64_t(len(pk.G1.K))
	d.g1_z, d.n_z = ptrOf(pk.G1.Z), C.uint64_t(len(pk.G1.Z))
	d.g2_b, d.n_b2 = ptrOf(pk.G2.B), C.uint64_t(len(pk.G2.B))
	d.infinity_a, d.infinity_b = (*C.uint8_t)(ptrOf(infA)), (*C.uint8_t)(ptrOf(infB))
	d.n_wires = C.uint64_t(len(pk.InfinityA))
	d.k_index = (*C.uint32_t)(ptrOf(kIndex))
	d.shard_count = 1
	var h C.uint64_t
	if rc := C.g16_pk_upload(ctx, &d, &h); rc != 0 {
		err := lastError(ctx, "g16_pk_upload")
		C.g16_ctx_destroy(ctx)
		return err
	}
	// the key becomes visible to Prove only once everything it needs is resident: a failed commitment-key upload
	// leaves pk.deviceInfo nil (the next Prove retries) instead of a half-initialised device state
	info := &deviceInfo{ctx: ctx, handle: uint64(h)}
	if err := info.uploadCommitmentKeys(&pk.ProvingKey); err != nil {
		C.g16_ctx_destroy(ctx)
		return err
	}
	pk.deviceInfo = info
	return nil
}

// newContext creates the library context over every visible GPU the library can group (1, 2, 4 or 8: the largest
// power of two not above the device count; G16_DEVICES=n caps it).  With more than one GPU the key is placed across
// them by g16_pk_upload and a single g16_prove_bn254 call uses them all (include/g16b200.h, g16_ctx_create).
func newContext() (ctxHandle, error) {
	n := int(C.g16_device_count())
	if n < 1 {
		return nil, fmt.Errorf("g16b200: no CUDA device (this backend has no CPU fallback)")
	}
	if v, err := strconv.Atoi(os.Getenv("G16_DEVICES")); err == nil && v >= 1 && v < n {
		n = v
	}
	world := 1
	for world*2 <= n && world < 8 {
		world *= 2
	}
	ids := make([]C.int, world)
	for i := range ids {
		ids[i] = C.int(i)
	}
	var ctx ctxHandle
	if rc := C.g16_ctx_create(&ctx, C.G16_CURVE_BN254, &ids[0], C.int(world)); rc != 0 {
		return nil, fmt.Errorf("g16_ctx_create over %d GPU(s): status %d", world, int(rc))
	}
	return ctx, nil
}

// BSB22 commitment keys (pedersen.ProvingKey{Basis, BasisExpSigma}) stay resident next to the key.
func (d *deviceInfo) uploadCommitmentKeys(pk *groth16_bn254.ProvingKey) error {
	for i := range pk.CommitmentKeys {
		ck := &pk.CommitmentKeys[i]
		var kh C.uint64_t
		if rc := C.g16_commitment_key_upload(d.ctx, ptrOf(ck.Basis), ptrOf(ck.BasisExpSigma), C.uint64_t(len(ck.Basis)), &kh); rc != 0 {
			return lastError(d.ctx, "g16_commitment_key_upload")
		}
		d.commitKeys = append(d.commitKeys, uint64(kh))
	}
	return nil
}

// Prove follows the reference prover: CPU fallback unless the accelerator option is set, solve on the CPU,
// device section (computeH + five MSMs) through one C call, O(1) epilogue on the CPU.
func Prove(r1cs *cs.R1CS, pk *ProvingKey, fullWitness witness.Witness, opts ...backend.ProverOption) (*groth16_bn254.Proof, error) {
	opt, err := backend.NewProverConfig(opts...)
	if err != nil {
		return nil, fmt.Errorf("new prover config: %w", err)
	}
	if opt.Accelerator != "icicle" {
		return groth16_bn254.Prove(r1cs, &pk.ProvingKey, fullWitness, opts...)
	}
	log := logger.Logger().With().Str("curve", r1cs.CurveID().String()).Str("acceleration", "g16b200").Int("nbConstraints", r1cs.GetNbConstraints()).Str("backend", "groth16").Logger()
	if opt.HashToFieldFn == nil {
		opt.HashToFieldFn = hash_to_field.New([]byte(constraint.CommitmentDst))
	}
	commitmentInfo := r1cs.CommitmentInfo.(constraint.Groth16Commitments)
	if err := pk.setupDevicePointers(r1cs.GetNbPublicVariables(), commitmentInfo); err != nil {
		return nil, err
	}
	proof := &groth16_bn254.Proof{Commitments: make([]curve.G1Affine, len(commitmentInfo))}
	poks := make([]curve.G1Affine, len(commitmentInfo))

	// the bsb22 hint (prove.go:81-109) with Commit moved to the device; the proof of knowledge (prove.go:120-127) is the
	// same values against BasisExpSigma, so g16_commit returns it from the same upload
	solverOpts := opt.SolverOpts[:len(opt.SolverOpts):len(opt.SolverOpts)]
	bsb22ID := solver.GetHintID(fcs.Bsb22CommitmentComputePlaceholder)
	solverOpts = append(solverOpts, solver.OverrideHint(bsb22ID, func(_ *big.Int, in []*big.Int, out []*big.Int) error {
		i := int(in[0].Int64())
		in = in[1:]
		hashed := in[:len(commitmentInfo[i].PublicAndCommitmentCommitted)]
		committed := in[len(hashed):]
		values := make([]fr.Element, len(committed))
		for j, inJ := range committed {
			values[j].SetBigInt(inJ)
		}
		var pin runtime.Pinner
		if len(values) != 0 {
			pin.Pin(&values[0])
		}
		// one call in flight per library context: hints may run on several solver goroutines, and other goroutines may
		// be proving with the same key
		pk.mu.Lock()
		rc := C.g16_commit(pk.ctx, C.uint64_t(pk.commitKeys[i]), ptrOf(values), C.uint64_t(len(values)),
			unsafe.Pointer(&proof.Commitments[i]), unsafe.Pointer(&poks[i]))
		pk.mu.Unlock()
		pin.Unpin()
		if rc != 0 {
			return lastError(pk.ctx, "g16_commit")
		}
		opt.HashToFieldFn.Write(constraint.SerializeCommitment(proof.Commitments[i].Marshal(), hashed, (fr.Bits-1)/8+1))
		hashBts := opt.HashToFieldFn.Sum(nil)
		opt.HashToFieldFn.Reset()
		nbBuf := fr.Bytes
		if opt.HashToFieldFn.Size() < fr.Bytes {
			nbBuf = opt.HashToFieldFn.Size()
		}
		var res fr.Element
		res.SetBytes(hashBts[:nbBuf])
		res.BigInt(out[0])
		return nil
	}))

	_solution, err := r1cs.Solve(fullWitness, solverOpts...)
	if err != nil {
		return nil, err
	}
	solution := _solution.(*cs.R1CSSolution)
	start := time.Now()

	// fold the proofs of knowledge (prove.go:125-135)
	if len(commitmentInfo) != 0 {
		commitmentsSerialized := make([]byte, fr.Bytes*len(commitmentInfo))
		for i := range commitmentInfo {
			copy(commitmentsSerialized[fr.Bytes*i:], solution.W[commitmentInfo[i].CommitmentIndex].Marshal())
		}
		challenge, err := fr.Hash(commitmentsSerialized, []byte("G16-BSB22"), 1)
		if err != nil {
			return nil, err
		}
		if _, err = proof.CommitmentPok.Fold(poks, challenge[0], ecc.MultiExpConfig{NbTasks: 1}); err != nil {
			return nil, err
		}
	}

	// uint64 backing store: the block is reinterpreted as G1Affine / G2Affine below, which need 8-byte alignment
	var out [C.G16_PROVE_OUT_BYTES / 8]uint64
	var stats C.g16_stats
	var pin runtime.Pinner
	pin.Pin(&solution.W[0])
	pin.Pin(&solution.A[0])
	pin.Pin(&solution.B[0])
	pin.Pin(&solution.C[0])
	// The slices are ordinary (pageable) Go heap memory: the library stages them through its pinned ring with its own
	// copier threads ("stage_threads"), W first so that the MSMs over W run while a, b, c are still being staged.
	pk.mu.Lock()
	rc := C.g16_prove_bn254(pk.ctx, C.uint64_t(pk.handle), unsafe.Pointer(&solution.W[0]), C.uint64_t(len(solution.W)),
		unsafe.Pointer(&solution.A[0]), unsafe.Pointer(&solution.B[0]), unsafe.Pointer(&solution.C[0]),
		C.uint64_t(len(solution.A)), unsafe.Pointer(&out[0]), &stats)
	pk.mu.Unlock()
	pin.Unpin()
	if rc != 0 {
		return nil, lastError(pk.ctx, "g16_prove_bn254")
	}
	// the five raw MSM results are gnark memory images
	arMsm := *(*curve.G1Affine)(unsafe.Pointer(&out[0]))
	bs1Msm := *(*curve.G1Affine)(unsafe.Pointer(&out[8]))
	krsK := *(*curve.G1Affine)(unsafe.Pointer(&out[16]))
	krsHZ := *(*curve.G1Affine)(unsafe.Pointer(&out[24]))
	bs2Msm := *(*curve.G2Affine)(unsafe.Pointer(&out[32]))

	// epilogue: r, s, kr = -r*s; deltas = [r, s, kr] * delta
	var _r, _s, _kr fr.Element
	if _, err := _r.SetRandom(); err != nil {
		return nil, err
	}
	if _, err := _s.SetRandom(); err != nil {
		return nil, err
	}
	_kr.Mul(&_r, &_s).Neg(&_kr)
	var r, s big.Int
	_r.BigInt(&r)
	_s.BigInt(&s)
	deltas := curve.BatchScalarMultiplicationG1(&pk.G1.Delta, []fr.Element{_r, _s, _kr})

	var ar, bs1, krs, p1 curve.G1Jac
	ar.FromAffine(&arMsm)
	ar.AddMixed(&pk.G1.Alpha)
	ar.AddMixed(&deltas[0])
	proof.Ar.FromJacobian(&ar)
	bs1.FromAffine(&bs1Msm)
	bs1.AddMixed(&pk.G1.Beta)
	bs1.AddMixed(&deltas[1])
	krs.FromAffine(&krsK)
	krs.AddMixed(&deltas[2])
	krs.AddMixed(&krsHZ)
	p1.ScalarMultiplication(&ar, &s)
	krs.AddAssign(&p1)
	p1.ScalarMultiplication(&bs1, &r)
	krs.AddAssign(&p1)
	proof.Krs.FromJacobian(&krs)

	var bs, deltaS curve.G2Jac
	bs.FromAffine(&bs2Msm)
	deltaS.FromAffine(&pk.G2.Delta)
	deltaS.ScalarMultiplication(&deltaS, &s)
	bs.AddAssign(&deltaS)
	bs.AddMixed(&pk.G2.Beta)
	proof.Bs.FromJacobian(&bs)

	log.Debug().Dur("took", time.Since(start)).Float32("device_ms", float32(stats.ms_total)).Msg("prover done")
	return proof, nil
}
