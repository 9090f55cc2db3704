// Source-only: not compiled in this repository's build image (no Go toolchain). See INTEGRATION.md.
//go:build cgo && cuda

package engine

import (
	"encoding/binary"
	"fmt"

	"github.com/shizukutanaka/Otedama/internal/miner"
	"github.com/shizukutanaka/Otedama/internal/poolproto"
	"github.com/shizukutanaka/Otedama/internal/poolproto/stratumv1"
)

// Stratum V1 on GPUs (SURVEY §8f-1): the device rolls extranonce2 and rebuilds the merkle root itself, so
// the engine must hand it the coinbase pieces and send back the extranonce2 that was actually hashed.
//
// applyGPUJob replaces applyJob (run.go:1340-1364) for GPU workers. What changes against the reference:
//   - the job id stays the pool's string (no Sscanf("%d") into a uint32: hex ids would be refused or
//     truncated) and travels in GPUWorkExt.JobIDStr; it comes back on every share;
//   - Version and PrevHash ARE copied into the header (the reference leaves them zero, run.go:1349-1359);
//     PrevHash is un-swapped from the notify wire order (eight 4-byte words, each reversed);
//   - coinb1 / coinb2 / merkle_branch / extranonce1 / extranonce2_size come from the session's
//     V1JobExt side table (stratumv1/notify_ext.go) instead of being dropped (parse.go:52-54);
//   - extranonce2 values are split over the G devices (e -> device e mod G).
func applyGPUJob(workers []*miner.GPUWorker, job poolproto.Job, ext stratumv1.V1JobExt, chanID uint32,
	difficulty float64) error {
	target, err := v1JobTarget(job.NBits, difficulty)
	if err != nil {
		return fmt.Errorf("engine: bad target for job %q: %w", job.JobID, err)
	}
	var prev [32]byte
	for w := 0; w < 8; w++ { // wire order -> header order
		for b := 0; b < 4; b++ {
			prev[4*w+b] = job.PrevHash[4*w+3-b]
		}
	}
	work := &miner.Work{ChannelID: chanID, NBits: job.NBits, Target: target,
		Header: miner.Header{Version: job.Version, PrevHash: prev, Time: job.NTime, Bits: job.NBits}}
	count := uint64(1) << 40 // every value the field can hold, capped (x 2^32 nonces each)
	if ext.ExtraNonce2Size < 5 {
		count = uint64(1) << (8 * uint(ext.ExtraNonce2Size))
	}
	return setPartitionedWork(workers, work, miner.GPUWorkExt{
		Coinb1: nonNil(ext.Coinb1), ExtraNonce1: ext.ExtraNonce1, Coinb2: ext.Coinb2,
		MerkleBranches: ext.MerkleBranch, ExtraNonce2Size: uint32(ext.ExtraNonce2Size),
		ExtraNonce2Count: count, ExtraNonce2Stride: 1, JobIDStr: job.JobID,
		VersionMask: ext.VersionMask, // 0 unless mining.configure negotiated version-rolling (BIP-310)
	})
}

func nonNil(b []byte) []byte {
	if b == nil {
		return []byte{}
	}
	return b
}

// gpuShareSubmission turns a GPU share into what session.Submit sends (stratumv1.go:330-366). With
// ExtraNonce set, Submit hex-encodes it instead of padding zeros (stratumv1.go:336-340); the bytes are the
// little-endian counter of extranonce2_size bytes that the device placed in the coinbase — the same rule
// otd_share_to_v1_submit applies for C callers.
func gpuShareSubmission(s miner.GPUShare) poolproto.ShareSubmission {
	var le [8]byte
	binary.LittleEndian.PutUint64(le[:], s.ExtraNonce2)
	en2 := make([]byte, s.ExtraNonce2Size)
	copy(en2, le[:]) // sizes above 8 keep zero bytes past the 8th, like the device
	jobID := s.JobIDStr
	if jobID == "" {
		jobID = fmt.Sprintf("%d", s.JobID)
	}
	return poolproto.ShareSubmission{JobID: jobID, Nonce: s.Nonce, NTime: s.NTime, Version: s.Version, ExtraNonce: en2}
}
