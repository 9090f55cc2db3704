// Source-only: not compiled in this repository's build image (no Go toolchain). See INTEGRATION.md.

package stratumv1

import (
	"encoding/hex"
	"encoding/json"
	"fmt"
	"sync"
)

// V1JobExt carries what parseNotify drops (parse.go:52-54: coinb1, coinb2, merkle_branch) and what the
// session keeps to itself (stratumv1.go:113-115: extranonce1, extranonce2_size): the inputs a device needs
// to roll extranonce2 and rebuild the merkle root on its own. Kept beside poolproto.Job, keyed by the
// job id, so that poolproto.Job and its consumers stay as they are.
type V1JobExt struct {
	Coinb1, Coinb2  []byte
	MerkleBranch    [][32]byte
	ExtraNonce1     []byte
	ExtraNonce2Size int
	VersionMask     uint32 // BIP-310 mask from mining.configure / mining.set_version_mask; 0 = no rolling
}

// parseNotifyExt decodes params 2..4 of mining.notify
// ([job_id, prevhash, coinb1, coinb2, merkle_branch, version, nbits, ntime, clean_jobs]).
func parseNotifyExt(raw json.RawMessage) (V1JobExt, error) {
	var p []json.RawMessage
	if err := json.Unmarshal(raw, &p); err != nil {
		return V1JobExt{}, err
	}
	if len(p) < 9 {
		return V1JobExt{}, fmt.Errorf("notify: expected 9 params, got %d", len(p))
	}
	var c1, c2 string
	var branch []string
	if err := json.Unmarshal(p[2], &c1); err != nil {
		return V1JobExt{}, err
	}
	if err := json.Unmarshal(p[3], &c2); err != nil {
		return V1JobExt{}, err
	}
	if err := json.Unmarshal(p[4], &branch); err != nil {
		return V1JobExt{}, err
	}
	var ext V1JobExt
	var err error
	if ext.Coinb1, err = hex.DecodeString(c1); err != nil {
		return V1JobExt{}, fmt.Errorf("notify: coinb1: %w", err)
	}
	if ext.Coinb2, err = hex.DecodeString(c2); err != nil {
		return V1JobExt{}, fmt.Errorf("notify: coinb2: %w", err)
	}
	ext.MerkleBranch = make([][32]byte, len(branch))
	for i, h := range branch {
		b, err := hex.DecodeString(h)
		if err != nil || len(b) != 32 {
			return V1JobExt{}, fmt.Errorf("notify: merkle_branch[%d] is not 32 hex-encoded bytes", i)
		}
		copy(ext.MerkleBranch[i][:], b)
	}
	return ext, nil
}

// jobExtTable is the session's side table. The read loop stores an entry next to each job it queues:
//
//	case "mining.notify":                                   (stratumv1.go read loop)
//	    job, err := parseNotify(msg.Params)
//	    if ext, xerr := parseNotifyExt(msg.Params); xerr == nil {     // + these four lines
//	        ext.ExtraNonce1, _ = hex.DecodeString(s.extranonce1)
//	        ext.ExtraNonce2Size = s.extranonce2Size
//	        s.jobExt.put(job.JobID, ext, job.CleanJobs)
//	    }
//
// clean_jobs empties the table first (older jobs are stale); otherwise it is bounded.
type jobExtTable struct {
	mu    sync.Mutex
	byID  map[string]V1JobExt
	order []string
}

const jobExtKeep = 16

func (t *jobExtTable) put(id string, ext V1JobExt, clean bool) {
	t.mu.Lock()
	defer t.mu.Unlock()
	if clean || t.byID == nil {
		t.byID, t.order = map[string]V1JobExt{}, nil
	}
	if _, ok := t.byID[id]; !ok {
		t.order = append(t.order, id)
	}
	t.byID[id] = ext
	for len(t.order) > jobExtKeep {
		delete(t.byID, t.order[0])
		t.order = t.order[1:]
	}
}

func (t *jobExtTable) get(id string) (V1JobExt, bool) {
	t.mu.Lock()
	defer t.mu.Unlock()
	e, ok := t.byID[id]
	return e, ok
}

// JobExt is the accessor the engine uses (exposed on the session through a small interface,
// `interface{ JobExt(jobID string) (V1JobExt, bool) }`, type-asserted like PoolNoticeReceiver):
//
//	func (s *session) JobExt(jobID string) (V1JobExt, bool) { return s.jobExt.get(jobID) }
