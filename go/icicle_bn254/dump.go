//go:build icicle

package icicle_bn254

/*
#include "g16b200.h"
*/
import "C"

import (
	"encoding/binary"
	"fmt"
	"io"
	"os"
	"runtime"
	"unsafe"

	curve "github.com/consensys/gnark-crypto/ecc/bn254"
	"github.com/consensys/gnark-crypto/ecc/bn254/fr/pedersen"
	gcunsafe "github.com/consensys/gnark-crypto/utils/unsafe"
)

// ReadDumpToDevice is ProvingKey.ReadDump (backend/groth16/bn254/marshal.go:458-550) with the large point slices
// streamed file -> pinned staging -> HBM by libg16b200 instead of being read into Go slices: a WriteDump image
// holds them as raw memory (gnark-crypto utils/unsafe.WriteSlice), which is already the device format.  The header is decoded by
// gnark-crypto exactly as ReadDump does; pk.G1.A/B/Z/K and pk.G2.B stay empty on the host.  kIndex is nil for
// circuits without commitments (see setupDevicePointers).
func (pk *ProvingKey) ReadDumpToDevice(f *os.File, kIndex []uint32) error {
	if err := gcunsafe.ReadMarker(f); err != nil {
		return fmt.Errorf("read marker: %w", err)
	}
	if _, err := pk.Domain.ReadFrom(f); err != nil {
		return fmt.Errorf("read domain: %w", err)
	}
	dec := curve.NewDecoder(f, curve.NoSubgroupChecks())
	var nbWires uint64
	var nbCommitments uint32
	for i, v := range []interface{}{&pk.G1.Alpha, &pk.G1.Beta, &pk.G1.Delta, &pk.G2.Beta, &pk.G2.Delta, &nbWires, &pk.NbInfinityA, &pk.NbInfinityB} {
		if err := dec.Decode(v); err != nil {
			return fmt.Errorf("read field %d: %w", i, err)
		}
	}
	pk.InfinityA = make([]bool, nbWires)
	pk.InfinityB = make([]bool, nbWires)
	for _, v := range []interface{}{&pk.InfinityA, &pk.InfinityB, &nbCommitments} {
		if err := dec.Decode(v); err != nil {
			return err
		}
	}
	// slice headers only: u64 length, then skip the raw elements
	span := func(elemBytes int64) (off, n uint64, err error) {
		if err = binary.Read(f, binary.LittleEndian, &n); err != nil {
			return
		}
		pos, err := f.Seek(0, io.SeekCurrent)
		if err != nil {
			return
		}
		_, err = f.Seek(int64(n)*elemBytes, io.SeekCurrent)
		return uint64(pos), n, err
	}
	var d C.g16_pk_file_desc
	var err error
	var off, n uint64
	if off, n, err = span(64); err != nil {
		return fmt.Errorf("read G1.A: %w", err)
	}
	d.off_g1_a, d.key.n_a = C.uint64_t(off), C.uint64_t(n)
	if off, n, err = span(64); err != nil {
		return fmt.Errorf("read G1.B: %w", err)
	}
	d.off_g1_b, d.key.n_b = C.uint64_t(off), C.uint64_t(n)
	if off, n, err = span(64); err != nil {
		return fmt.Errorf("read G1.Z: %w", err)
	}
	d.off_g1_z, d.key.n_z = C.uint64_t(off), C.uint64_t(n)
	if off, n, err = span(64); err != nil {
		return fmt.Errorf("read G1.K: %w", err)
	}
	d.off_g1_k, d.key.n_k = C.uint64_t(off), C.uint64_t(n)
	if off, n, err = span(128); err != nil {
		return fmt.Errorf("read G2.B: %w", err)
	}
	d.off_g2_b, d.key.n_b2 = C.uint64_t(off), C.uint64_t(n)
	// commitment keys are small: read them the ordinary way (marshal.go:530-543)
	for i := 0; i < int(nbCommitments); i++ {
		var ck pedersen.ProvingKey
		if ck.Basis, _, err = gcunsafe.ReadSlice[[]curve.G1Affine](f); err != nil {
			return fmt.Errorf("read commitment basis %d: %w", i, err)
		}
		if ck.BasisExpSigma, _, err = gcunsafe.ReadSlice[[]curve.G1Affine](f); err != nil {
			return fmt.Errorf("read commitment basisExpSigma %d: %w", i, err)
		}
		pk.CommitmentKeys = append(pk.CommitmentKeys, ck)
	}

	ctx, err := newContext()
	if err != nil {
		return err
	}
	infA, infB := boolBytes(pk.InfinityA), boolBytes(pk.InfinityB)
	var pin runtime.Pinner
	defer pin.Unpin()
	for _, p := range []unsafe.Pointer{ptrOf(infA), ptrOf(infB), ptrOf(kIndex)} {
		if p != nil {
			pin.Pin(p)
		}
	}
	d.key.domain_n = C.uint64_t(pk.Domain.Cardinality)
	d.key.infinity_a, d.key.infinity_b = (*C.uint8_t)(ptrOf(infA)), (*C.uint8_t)(ptrOf(infB))
	d.key.n_wires = C.uint64_t(nbWires)
	d.key.k_index = (*C.uint32_t)(ptrOf(kIndex))
	d.key.shard_count = 1
	d.fd = C.int(f.Fd())
	var h C.uint64_t
	if rc := C.g16_pk_upload_from_file(ctx, &d, &h); rc != 0 {
		err := lastError(ctx, "g16_pk_upload_from_file")
		C.g16_ctx_destroy(ctx)
		return err
	}
	info := &deviceInfo{ctx: ctx, handle: uint64(h)}
	if err := info.uploadCommitmentKeys(&pk.ProvingKey); err != nil {
		C.g16_ctx_destroy(ctx)
		return err
	}
	pk.deviceInfo = info
	return nil
}
