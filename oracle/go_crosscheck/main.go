// go_crosscheck pins the repository's oracle against the real Go standard library.
//
// Test infrastructure only (like everything under oracle/).  Stdlib only: no module download is needed.
// It restates, with the reference's own calls, the per-token body of the two hot loops
//
//	adapter  json.Unmarshal(msg.Data, &token)        src/sse-adapter/sse_handler.go:136-140
//	         startSequence filter                    src/sse-adapter/sse_handler.go:143
//	         json.Marshal(token) + three fmt.Fprintf src/sse-adapter/sse_handler.go:211-214
//	bridge   json.Unmarshal / subject / json.Marshal src/redis-nats-bridge/main.go:153-161
//
// replays the committed vectors (tests/golden/adapter_vectors.json.gz, written by tests/golden/make_golden.py,
// and tests/golden/hand_vectors.json, written by tests/golden/export_hand_vectors.py) and compares status and
// bytes with what the vectors expect.  Exit status 0 and a line "OK <n> vectors, go<version>" mean the
// oracle, the Python model and the CUDA path (which all reproduce these vectors) agree with real Go.
//
// NOT COMPILED in the build image of this repository (no Go toolchain there or on the GPU box):
// tests/golden/run_go_crosscheck.sh builds and runs it wherever `go version` works.
package main

import (
	"bytes"
	"compress/gzip"
	"encoding/hex"
	"encoding/json"
	"fmt"
	"io"
	"os"
	"runtime"
	"strconv"
	"strings"
)

// TokenMessage is the reference's wire schema, field for field (src/sse-adapter/sse_handler.go:46-52).
type TokenMessage struct {
	ConversationID string `json:"conversation_id"`
	Token          string `json:"token"`
	Sequence       int64  `json:"sequence"`
	Done           bool   `json:"done"`
	Timestamp      int64  `json:"timestamp"`
}

// vector is one entry of adapter_vectors.json.gz / hand_vectors.json (hex strings).
type vector struct {
	Name     string `json:"name"`
	In       string `json:"in"`
	Start    int64  `json:"start"`
	Status   uint32 `json:"status"`
	Frame    string `json:"frame"`
	Frame122 string `json:"frame122"` // present when Go >= 1.22 spells the BS and FF controls differently
	BStatus  uint32 `json:"bstatus"`
	Subject  string `json:"subject"`
	JSON     string `json:"json"`
	JSON122  string `json:"json122"`
	NoBridge  bool  `json:"nobridge"`  // hand vectors that only state the adapter side
	NoAdapter bool  `json:"noadapter"` // hand vectors that only state the bridge side
}

const (
	stOK, stErrSyntax, stErrType, stFiltered = 0, 1, 2, 3
	stDone                                   = 0x100
)

// decode is json.Unmarshal as both services call it; the status code separates the two error kinds the
// library reports (both make the reference log and drop the message).
func decode(payload []byte) (TokenMessage, uint32) {
	var token TokenMessage
	if err := json.Unmarshal(payload, &token); err != nil {
		if _, ok := err.(*json.SyntaxError); ok {
			return token, stErrSyntax
		}
		return token, stErrType // *json.UnmarshalTypeError
	}
	return token, stOK
}

// adapter: sse_handler.go:136-145 (callback) + :211-214 (frame)
func adapter(payload []byte, startSequence int64) ([]byte, uint32) {
	token, st := decode(payload)
	if st != stOK {
		return nil, st
	}
	if startSequence > 0 && token.Sequence <= startSequence {
		return nil, stFiltered
	}
	var w bytes.Buffer
	data, _ := json.Marshal(token)
	fmt.Fprintf(&w, "event: token\n")
	fmt.Fprintf(&w, "id: %d\n", token.Sequence)
	fmt.Fprintf(&w, "data: %s\n\n", data)
	if token.Done {
		st |= stDone
	}
	return w.Bytes(), st
}

// bridge: redis-nats-bridge/main.go:153-161
func bridge(payload []byte) (string, []byte, uint32) {
	token, st := decode(payload)
	if st != stOK {
		return "", nil, st
	}
	subject := fmt.Sprintf("chat.%s.tokens", token.ConversationID)
	data, _ := json.Marshal(token)
	if token.Done {
		st |= stDone
	}
	return subject, data, st
}

func load(path string) ([]vector, error) {
	f, err := os.Open(path)
	if err != nil {
		return nil, err
	}
	defer f.Close()
	var r io.Reader = f
	if strings.HasSuffix(path, ".gz") {
		z, err := gzip.NewReader(f)
		if err != nil {
			return nil, err
		}
		defer z.Close()
		r = z
	}
	var v []vector
	err = json.NewDecoder(r).Decode(&v)
	return v, err
}

// go122 reports whether this toolchain writes the BS and FF controls as backslash-b / backslash-f
// (Go >= 1.22) instead of the six-character u0008 / u000c escapes of Go <= 1.21.
func go122() bool {
	v := strings.TrimPrefix(runtime.Version(), "go")
	parts := strings.SplitN(v, ".", 3)
	if len(parts) < 2 {
		return true // development toolchains are newer than 1.22
	}
	major, err1 := strconv.Atoi(parts[0])
	minorDigits := parts[1]
	for i, c := range minorDigits {
		if c < '0' || c > '9' {
			minorDigits = minorDigits[:i]
			break
		}
	}
	minor, err2 := strconv.Atoi(minorDigits)
	if err1 != nil || err2 != nil {
		return true
	}
	return major > 1 || minor >= 22
}

func main() {
	if len(os.Args) < 2 {
		fmt.Fprintln(os.Stderr, "usage: go_crosscheck <vectors.json[.gz]> ...")
		os.Exit(2)
	}
	newSpelling := go122()
	total, bad := 0, 0
	fail := func(name, what, got, want string) {
		bad++
		if bad <= 40 {
			fmt.Printf("MISMATCH %s %s\n  go   %s\n  want %s\n", name, what, got, want)
		}
	}
	for _, path := range os.Args[1:] {
		vecs, err := load(path)
		if err != nil {
			fmt.Fprintln(os.Stderr, path+":", err)
			os.Exit(2)
		}
		for _, c := range vecs {
			total++
			in, err := hex.DecodeString(c.In)
			if err != nil {
				fmt.Fprintln(os.Stderr, c.Name+": bad hex input")
				os.Exit(2)
			}
			wantFrame, wantJSON := c.Frame, c.JSON
			if newSpelling && c.Frame122 != "" {
				wantFrame = c.Frame122
			}
			if newSpelling && c.JSON122 != "" {
				wantJSON = c.JSON122
			} else if newSpelling && c.Frame122 != "" && c.JSON != "" {
				// the fixtures keep the Go >= 1.22 spelling only as a frame: its data line is the JSON
				if raw, err := hex.DecodeString(c.Frame122); err == nil {
					if k := bytes.Index(raw, []byte("\ndata: ")); k >= 0 && len(raw) >= k+7+2 {
						wantJSON = hex.EncodeToString(raw[k+7 : len(raw)-2])
					}
				}
			}
			if !c.NoAdapter {
				frame, st := adapter(in, c.Start)
				if st != c.Status {
					fail(c.Name, "adapter status", fmt.Sprint(st), fmt.Sprint(c.Status))
				}
				if hex.EncodeToString(frame) != wantFrame {
					fail(c.Name, "frame", hex.EncodeToString(frame), wantFrame)
				}
			}
			if c.NoBridge {
				continue
			}
			subject, data, bst := bridge(in)
			if bst != c.BStatus {
				fail(c.Name, "bridge status", fmt.Sprint(bst), fmt.Sprint(c.BStatus))
			}
			if hex.EncodeToString([]byte(subject)) != c.Subject {
				fail(c.Name, "subject", hex.EncodeToString([]byte(subject)), c.Subject)
			}
			if hex.EncodeToString(data) != wantJSON {
				fail(c.Name, "bridge json", hex.EncodeToString(data), wantJSON)
			}
		}
	}
	if bad != 0 {
		fmt.Printf("FAILED %d mismatches in %d vectors, %s\n", bad, total, runtime.Version())
		os.Exit(1)
	}
	fmt.Printf("OK %d vectors, %s\n", total, runtime.Version())
}
