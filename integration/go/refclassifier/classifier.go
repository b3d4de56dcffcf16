// Package refclassifier is the plain-Go reference classifier BASELINE.json's north_star names: the per-packet
// match-action lookup that the ServiceFunctionChain CR and the vendor plugin's bridge-port / network-function
// programming ask a DPU to perform, written against the repository's own types.
//
// UNCOMPILED: no Go toolchain exists in the build image or on the GPU box ("unavailable on this box", BASELINE.md).
// The executable definition of the same semantics, and the CPU baseline that is actually timed, is the C++ oracle
// (oracle/sfc_oracle.cpp); this file restates it in Go so that a maintainer with a toolchain can time it beside the
// GPU path and diff its answers (same stage order, same result word).  Semantics: SURVEY.md §8a.
package refclassifier

import (
	"bytes"
	"encoding/binary"
	"sort"

	configv1 "github.com/openshift/dpu-operator/api/v1"
)

// Result word: action:8 | next_nf:8 | out_port:16 | stage:4 | index:28 (include/b200sfc.h SFC_RESULT).
const (
	ActToCP, ActForward, ActDrop                                          = 0, 1, 2
	StageMiss, StageACL, StageFlow, StageL2Exact, StageL2Wild, StageLPM4, StageLPM6 = 0, 1, 2, 3, 4, 5, 6
	NFNone                                                                = 0xFF
)

func result(action, nf uint8, port uint16, stage uint8, index uint32) uint64 {
	return uint64(action)<<56 | uint64(nf)<<48 | uint64(port)<<32 | uint64(stage)<<28 | uint64(index&0x0FFFFFFF)
}

// L2Rule is one `ovs-ofctl add-flow` of the Marvell backend (marvell/ovs-dp/ovsdp.go:113-142).
type L2Rule struct {
	InPort, OutPort uint16
	DstMAC          []byte // nil: the in_port wildcard (priority 10)
	Priority        uint32 // 10 / 100 (hairpin) / 32768 (OpenFlow default)
	ID              uint32
}

type FlowKey struct {
	IPVer, Proto uint8
	VLAN         uint16
	Src, Dst     [16]byte
	SPort, DPort uint16
}
type FlowAction struct {
	Action  uint8
	OutPort uint16
	ID      uint32
}
type ACLRule struct {
	IPVer                                          uint8 // 0 = any
	InPort, InPortMask, VLAN, VLANMask             uint16
	Proto, ProtoMask                               uint8
	SPort, SPortMask, DPort, DPortMask             uint16
	Src, SrcMask, Dst, DstMask                     [16]byte
	Deny                                           bool
	Priority, ID                                   uint32
}
type Prefix struct {
	Addr    [16]byte
	Len     int
	OutPort uint16
	ID      uint32
}

type Classifier struct {
	NFIn, NFOut []uint16 // the chain: ports of NetworkFunction i (ServiceFunctionChainSpec.NetworkFunctions order)
	L2          []L2Rule
	Flows       map[FlowKey]FlowAction
	ACL         []ACLRule
	LPM4, LPM6  []Prefix
}

// VF is a host VF as the host-side daemon reports it (internal/daemon/hostsidemanager.go:48-74).
type VF struct {
	Port uint16
	MAC  []byte
}

// FromChain compiles the steering set the Marvell VSP programs (marvell/main.go:382-403, 519-560), generalised to
// the N network functions of the CR: VF -> NF[0].in; NF[i].out -> NF[i+1].in; last out <-> wire; towards a VF by dl_dst.
// nfPorts[i] = {ingress, egress} port ids of sfc.Spec.NetworkFunctions[i] (two per NF: sfc-reconciler/sfc.go:39,55).
func FromChain(sfc *configv1.ServiceFunctionChain, nfPorts [][2]uint16, vfs []VF, wire uint16) *Classifier {
	c := &Classifier{Flows: map[FlowKey]FlowAction{}}
	n := len(sfc.Spec.NetworkFunctions)
	for i := 0; i < n; i++ {
		c.NFIn = append(c.NFIn, nfPorts[i][0])
		c.NFOut = append(c.NFOut, nfPorts[i][1])
	}
	add := func(in, out uint16, mac []byte) {
		prio := uint32(10)
		if mac != nil {
			prio = 32768
			if in == out {
				prio = 100
			}
		}
		c.L2 = append(c.L2, L2Rule{in, out, mac, prio, uint32(len(c.L2))})
	}
	last := c.NFOut[n-1]
	for _, vf := range vfs {
		add(vf.Port, c.NFIn[0], nil)
		add(c.NFIn[0], vf.Port, vf.MAC)
		add(last, last, vf.MAC)
		for i := 1; i < n; i++ {
			add(c.NFIn[i], c.NFOut[i-1], vf.MAC)
		}
	}
	for i := 0; i+1 < n; i++ {
		add(c.NFOut[i], c.NFIn[i+1], nil)
	}
	add(last, wire, nil)
	add(wire, last, nil)
	return c
}

func (c *Classifier) nfOf(port uint16) uint8 {
	for i := range c.NFIn {
		if c.NFIn[i] == port || c.NFOut[i] == port {
			return uint8(i)
		}
	}
	return NFNone
}

type parsed struct {
	dmac         []byte
	vlan         uint16
	ipVer, proto uint8
	src, dst     [16]byte
	sport, dport uint16
}

// parse reads the 64-byte record: one optional 802.1Q tag, IPv4 (IHL honoured) or IPv6 without extension headers,
// ports for TCP / UDP / SCTP of unfragmented packets; big-endian as on the wire (dpu-cni/pkgs/sriovutils/packet.go:41-66).
func parse(b []byte) parsed {
	p := parsed{dmac: b[0:6]}
	off := 12
	et := binary.BigEndian.Uint16(b[off:])
	off += 2
	if et == 0x8100 {
		p.vlan = binary.BigEndian.Uint16(b[off:]) & 0x0FFF
		et = binary.BigEndian.Uint16(b[off+2:])
		off += 4
	}
	l4 := func(proto uint8) bool { return proto == 6 || proto == 17 || proto == 132 }
	switch et {
	case 0x0800:
		ip := b[off:]
		ihl := int(ip[0] & 0x0F)
		if ip[0]>>4 != 4 || ihl < 5 {
			return p
		}
		p.ipVer, p.proto = 4, ip[9]
		copy(p.src[:4], ip[12:16])
		copy(p.dst[:4], ip[16:20])
		at := off + 4*ihl
		if l4(p.proto) && binary.BigEndian.Uint16(ip[6:])&0x1FFF == 0 && at+4 <= 64 {
			p.sport, p.dport = binary.BigEndian.Uint16(b[at:]), binary.BigEndian.Uint16(b[at+2:])
		}
	case 0x86DD:
		ip := b[off:]
		if ip[0]>>4 != 6 {
			return p
		}
		p.ipVer, p.proto = 6, ip[6]
		copy(p.src[:], ip[8:24])
		copy(p.dst[:], ip[24:40])
		if at := off + 40; l4(p.proto) && at+4 <= 64 {
			p.sport, p.dport = binary.BigEndian.Uint16(b[at:]), binary.BigEndian.Uint16(b[at+2:])
		}
	}
	return p
}

func masked16(a, b, m [16]byte) bool {
	for i := range a {
		if (a[i]^b[i])&m[i] != 0 {
			return false
		}
	}
	return true
}

// Classify: ACL (highest priority, ties to the lowest id) -> exact 5-tuple -> L2 steering (OpenFlow priority) -> LPM -> miss.
func (c *Classifier) Classify(hdr []byte, inPort uint16) uint64 {
	k := parse(hdr)
	if k.ipVer != 0 {
		var best *ACLRule
		for i := range c.ACL {
			r := &c.ACL[i]
			if (r.IPVer != 0 && r.IPVer != k.ipVer) || (inPort^r.InPort)&r.InPortMask != 0 || (k.vlan^r.VLAN)&r.VLANMask != 0 ||
				(k.proto^r.Proto)&r.ProtoMask != 0 || (k.sport^r.SPort)&r.SPortMask != 0 || (k.dport^r.DPort)&r.DPortMask != 0 ||
				!masked16(k.src, r.Src, r.SrcMask) || !masked16(k.dst, r.Dst, r.DstMask) {
				continue
			}
			if best == nil || r.Priority > best.Priority || (r.Priority == best.Priority && r.ID < best.ID) {
				best = r
			}
		}
		if best != nil && best.Deny {
			return result(ActDrop, NFNone, 0, StageACL, best.ID)
		}
		if f, ok := c.Flows[FlowKey{k.ipVer, k.proto, k.vlan, k.src, k.dst, k.sport, k.dport}]; ok {
			nf := uint8(NFNone)
			if f.Action == ActForward {
				nf = c.nfOf(f.OutPort)
			}
			return result(f.Action, nf, f.OutPort, StageFlow, f.ID)
		}
	}
	var hit *L2Rule
	for i := range c.L2 {
		r := &c.L2[i]
		if r.InPort != inPort || (r.DstMAC != nil && !bytes.Equal(r.DstMAC, k.dmac)) {
			continue
		}
		if hit == nil || r.Priority > hit.Priority || (r.Priority == hit.Priority && r.ID < hit.ID) {
			hit = r
		}
	}
	if hit != nil {
		stage := uint8(StageL2Wild)
		if hit.DstMAC != nil {
			stage = StageL2Exact
		}
		return result(ActForward, c.nfOf(hit.OutPort), hit.OutPort, stage, hit.ID)
	}
	if k.ipVer != 0 {
		table, stage, bits := c.LPM4, uint8(StageLPM4), 32
		if k.ipVer == 6 {
			table, stage, bits = c.LPM6, StageLPM6, 128
		}
		var best *Prefix
		for i := range table {
			p := &table[i]
			if p.Len > bits || !prefixCovers(p, k.dst) {
				continue
			}
			if best == nil || p.Len > best.Len {
				best = p
			}
		}
		if best != nil {
			return result(ActForward, c.nfOf(best.OutPort), best.OutPort, stage, best.ID)
		}
	}
	return result(ActToCP, NFNone, 0, StageMiss, 0)
}

func prefixCovers(p *Prefix, a [16]byte) bool {
	full, rem := p.Len/8, p.Len%8
	if !bytes.Equal(p.Addr[:full], a[:full]) {
		return false
	}
	return rem == 0 || (p.Addr[full]^a[full])>>(8-rem) == 0
}

// SortLPM is a convenience for callers that want deterministic iteration order when dumping the tables.
func (c *Classifier) SortLPM() {
	for _, t := range [][]Prefix{c.LPM4, c.LPM6} {
		sort.Slice(t, func(i, j int) bool { return t[i].ID < t[j].ID })
	}
}
