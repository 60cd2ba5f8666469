// Package preprocessor keeps the exported surface of llm-message-queue/internal/preprocessor
// (preprocessor.go:24, 56, 171, 176, 187, 192, 253) on top of the B200 engine.  SOURCE ONLY: not compiled in this image.
//
// ProcessMessage keeps its signature for existing callers; the call the GPU wants is the additive ProcessMessages,
// which packs a batch into a pinned arena of the engine's ring and runs it in one submit.
package preprocessor

import (
	"fmt"
	"regexp"
	"strings"
	"time"

	"llm-message-queue/pkg/config"
	"llm-message-queue/pkg/models"

	"llm-message-queue/internal/triage"
)

type Preprocessor struct {
	Engine   *triage.Engine
	patterns map[models.Priority][]*regexp.Regexp // what GetKeywordPatterns returns; the engine holds its own compiled form
	users    map[string]models.Priority
}

var priorityByName = map[string]uint8{"realtime": 1, "high": 2, "normal": 3, "low": 4}

// NewPreprocessor ignores cfg exactly like the reference does (preprocessor.go:24-53) and installs the default dictionary.
func NewPreprocessor(cfg config.QueueConfig) *Preprocessor {
	e, err := triage.New(0, false, 0)
	if err != nil {
		panic(err) // no CPU fallback: without a B200 there is no preprocessor
	}
	return NewOnEngine(e)
}

func NewOnEngine(e *triage.Engine) *Preprocessor {
	p := &Preprocessor{Engine: e, patterns: map[models.Priority][]*regexp.Regexp{}, users: map[string]models.Priority{}}
	for _, w := range []string{"immediate", "emergency", "asap", "right now"} {
		p.patterns[models.PriorityRealtime] = append(p.patterns[models.PriorityRealtime], regexp.MustCompile("(?i)"+w))
	}
	for _, w := range []string{"urgent", "important", "priority", "critical", "soon"} {
		p.patterns[models.PriorityHigh] = append(p.patterns[models.PriorityHigh], regexp.MustCompile("(?i)"+w))
	}
	return p
}

// MetaOf is what ProcessMessage reads of a message (preprocessor.go:63-86, 120-138; queue_factory.go:217).
func (p *Preprocessor) MetaOf(msg *models.Message) triage.Meta {
	var f triage.MetaFields
	pr := int(msg.Priority)
	if pr < 0 || pr > 254 {
		pr = 255 // any other explicit priority: left untouched by ProcessMessage
	}
	f.Preset = uint8(pr)
	if s, ok := msg.Metadata["user_priority"].(string); ok {
		if c, ok := priorityByName[strings.ToLower(s)]; ok {
			f.UserPriority = c
		} else {
			f.UserPriority = 5 // TRIAGE_UP_UNRECOGNISED: the branch is still taken (:68-82)
		}
	}
	if up, ok := p.users[msg.UserID]; ok {
		f.UserDefault = uint8(up)
	}
	if s, ok := msg.Metadata["priority"].(string); ok {
		f.MetaPriority = priorityByName[strings.ToLower(s)]
	}
	if _, ok := msg.Metadata["vip_user"]; ok {
		f.Flags = 1 // TRIAGE_META_VIP
	}
	return triage.MakeMeta(f)
}

// Apply writes back what ProcessMessage writes (preprocessor.go:58-113).
func Apply(msg *models.Message, r *triage.Result, now time.Time) {
	if msg.Metadata == nil {
		msg.Metadata = make(map[string]interface{})
	}
	if triage.Flags(r)&1 == 0 { // not analysed: an explicit priority is respected (:63-65)
		return
	}
	pre := models.Priority(triage.QueueLevel(r)) // the pre-rule priority names the queue (:103-105)
	if pre < 1 || pre > 4 {
		pre = models.Priority(triage.Label(r))
	}
	msg.Priority = models.Priority(triage.Label(r))
	switch triage.Reason(r) {
	case 1:
		msg.Metadata["priority_reason"] = "user_override"
	case 2:
		msg.Metadata["priority_reason"] = "user_default"
	case 3:
		msg.Metadata["priority_reason"] = "content_keywords"
	}
	if triage.Flags(r)&4 != 0 {
		msg.Metadata["word_count"] = int(triage.Words(r))
		msg.Metadata["sentiment"] = [...]string{"negative", "neutral", "positive"}[triage.Sentiment(r)+1]
		if triage.Question(r) {
			msg.Metadata["contains_question"] = "true"
		} else {
			msg.Metadata["contains_question"] = "false"
		}
	}
	msg.Metadata["analyzed"] = true
	if msg.QueueName == "" {
		msg.QueueName = fmt.Sprint(pre)
	}
	if msg.CreatedAt.IsZero() {
		msg.CreatedAt = now
	}
	msg.UpdatedAt = now
}

// ProcessMessages is ProcessMessage for a batch: one arena, one submit.
func (p *Preprocessor) ProcessMessages(msgs []*models.Message) []*models.Message {
	if len(msgs) == 0 {
		return msgs
	}
	var total uint64
	for _, m := range msgs {
		total += uint64(len(m.Content))
	}
	a, err := p.Engine.Ring(len(msgs), total)
	if err != nil {
		panic(err)
	}
	for _, m := range msgs {
		a.Append(m.Content, p.MetaOf(m))
	}
	if err := p.Engine.Process(a); err != nil {
		panic(err)
	}
	now := time.Now()
	for i, m := range msgs {
		Apply(m, &a.Results[i], now)
	}
	return msgs
}

func (p *Preprocessor) ProcessMessage(msg *models.Message) *models.Message {
	return p.ProcessMessages([]*models.Message{msg})[0]
}

func (p *Preprocessor) SetUserPriority(userID string, priority models.Priority) {
	p.users[userID] = priority
	p.Engine.SetUserPriority(userID, int(priority))
}

// AddKeywordPattern returns the compile error of the pattern, or the engine's reason when the pattern is outside the
// RE2 subset it runs (anchors, \b, named groups, ... - include/triage.h).
func (p *Preprocessor) AddKeywordPattern(priority models.Priority, pattern string) error {
	re, err := regexp.Compile(pattern)
	if err != nil {
		return err
	}
	if err := p.Engine.AddPattern(int(priority), pattern); err != nil {
		return err
	}
	p.patterns[priority] = append(p.patterns[priority], re)
	return nil
}

func (p *Preprocessor) SetDefaultPriority(priority models.Priority) { p.Engine.SetDefaultPriority(int(priority)) }

func (p *Preprocessor) GetKeywordPatterns(priority models.Priority) []*regexp.Regexp {
	return p.patterns[priority]
}

// AnalyzeMessageContent (preprocessor.go:253-299): the same analysis with is_question as a bool, also for "".
func (p *Preprocessor) AnalyzeMessageContent(content string) map[string]interface{} {
	out := map[string]interface{}{"word_count": 0, "sentiment": "neutral", "is_question": false}
	if content == "" {
		return out
	}
	a, err := p.Engine.Ring(1, uint64(len(content)))
	if err != nil {
		panic(err)
	}
	a.Append(content, triage.MakeMeta(triage.MetaFields{Preset: 3}))
	if err := p.Engine.Process(a); err != nil {
		panic(err)
	}
	r := &a.Results[0]
	out["word_count"] = int(triage.Words(r))
	out["sentiment"] = [...]string{"negative", "neutral", "positive"}[triage.Sentiment(r)+1]
	out["is_question"] = triage.Question(r)
	return out
}
