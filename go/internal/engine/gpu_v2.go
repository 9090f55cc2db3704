// Source-only: not compiled in this repository's build image (no Go toolchain). See INTEGRATION.md.
//go:build cgo && cuda

package engine

import (
	"github.com/shizukutanaka/Otedama/internal/miner"
	"github.com/shizukutanaka/Otedama/internal/stratum"
)

// Stratum V2 version rolling (SURVEY §8f-2).
//
// Where the mask comes from: the Mining Protocol's SetupConnection.Flags. The reference sends Flags = 0
// (run.go:1172-1184, the field exists: stratum/handshake.go:33-43). Bit 2 of the client's flags is
// REQUIRES_VERSION_ROLLING; a server that answers SetupConnection.Success to it lets the device roll the
// BIP-320 general-purpose bits of NewMiningJob.Version, and SubmitSharesStandard.NVersion carries the
// rolled version back (run.go:929-939 already echoes Share.Version). Two edits in handshake():
//
//	sc := stratum.SetupConnection{ ..., Flags: sv2FlagRequiresVersionRolling, DeviceID: "gpu-cuda" }
//	... on MsgSetupConnectionSuccess: versionMask = bip320Mask        (on MsgSetupConnectionError with
//	    the unsupported-flags code: retry with Flags = 0 and versionMask = 0)
const (
	sv2FlagRequiresVersionRolling uint32 = 1 << 2
	bip320Mask                    uint32 = 0x1fffe000 // BIP-320: bits 13..28
)

// updateGPUWork is updateWork (run.go:1271-1297) for GPU workers: the same Work, plus — when the channel
// negotiated version rolling — the mask and the static partition of the 2^16 version indices over the G
// devices (version index k -> device k mod G; each device then searches versions x 2^32 nonces and goes
// idle when its slice is exhausted, instead of rescanning one header forever).
func updateGPUWork(workers []*miner.GPUWorker, job *stratum.NewMiningJob, chanID uint32, prevHash [32]byte,
	prevNBits uint32, ntime uint32, shareTarget miner.Hash, versionMask uint32) error {
	target := shareTarget
	if target == (miner.Hash{}) {
		t, err := miner.TargetFromNBits(prevNBits)
		if err != nil {
			return nil // the reference returns silently here too
		}
		target = t
	}
	w := &miner.Work{JobID: job.JobID, ChannelID: chanID, NBits: prevNBits, Target: target,
		Header: miner.Header{Version: job.Version, PrevHash: prevHash, MerkleRoot: job.MerkleRoot,
			Time: ntime, Bits: prevNBits}}
	if versionMask == 0 {
		for _, wr := range workers { // no rolling negotiated: the reference's behaviour, every device the same header
			wr.SetWork(w)
		}
		return nil
	}
	return setPartitionedWork(workers, w, miner.GPUWorkExt{VersionMask: versionMask})
}
