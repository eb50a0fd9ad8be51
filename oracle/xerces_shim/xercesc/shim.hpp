// TEST INFRASTRUCTURE (oracle build only) -- never linked into the product library.
//
// A minimal, header-only stand-in for the slice of the Xerces-C 3 API that the reference's
// loader touches (utils/xml/SUMOSAXReader.cpp, XMLSubSys.cpp, GenericSAXHandler.cpp,
// SUMOSAXAttributesImpl_Xerces.cpp, utils/options/OptionsIO.cpp, OptionsLoader.cpp,
// utils/handlers/TemplateHandler.cpp).  Xerces-C is a REQUIRED dependency of the reference
// (CMakeLists.txt:476) and is absent from this image; libexpat is present.  With this shim the
// unmodified reference sources compile into oracle/_ref/sumo, which is the parity oracle.
//
// Design: expat callbacks never call into the SAX handlers (they may throw C++ exceptions, which
// must not unwind through C frames).  Instead the callbacks queue events; the reader dispatches
// them from C++ after XML_Parse returns.  Progressive parsing (parseFirst/parseNext) delivers one
// queued event per parseNext call, which is what SUMORouteLoader::loadUntil and
// OptionsIO::getRoot rely on.  Validation is not implemented: run with --xml-validation never.
#pragma once

#include <expat.h>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <deque>
#include <fstream>
#include <memory>
#include <string>
#include <vector>

#define XERCES_CPP_NAMESPACE xercesc
#define XERCES_CPP_NAMESPACE_USE using namespace xercesc;
#define _XERCES_VERSION 30000   // < 30100 selects the plain XMLString::transcode paths in XMLSubSys.cpp:215

typedef char16_t XMLCh;
typedef unsigned char XMLByte;
typedef std::size_t XMLSize_t;
typedef std::uint64_t XMLFilePos;
typedef std::uint64_t XMLFileLoc;

namespace xercesc {

class MemoryManager {};

class XMLPlatformUtils {
public:
    static void Initialize() {}
    static void Terminate() {}
    static inline MemoryManager* fgMemoryManager = nullptr;
};

// ---------------------------------------------------------------------------------------------
// strings: XMLCh is UTF-16, the narrow side is UTF-8
// ---------------------------------------------------------------------------------------------
class XMLString {
public:
    static XMLSize_t stringLen(const XMLCh* s) {
        if (s == nullptr) {
            return 0;
        }
        XMLSize_t n = 0;
        while (s[n] != 0) {
            n++;
        }
        return n;
    }
    static bool equals(const XMLCh* a, const XMLCh* b) {
        if (a == nullptr || b == nullptr) {
            return a == b;
        }
        while (*a != 0 && *a == *b) {
            a++;
            b++;
        }
        return *a == *b;
    }
    static std::u16string toU16(const char* utf8, std::size_t len) {
        std::u16string out;
        out.reserve(len);
        std::size_t i = 0;
        while (i < len) {
            const unsigned char c = (unsigned char)utf8[i];
            std::uint32_t cp;
            int extra;
            if (c < 0x80) {
                cp = c;
                extra = 0;
            } else if ((c >> 5) == 0x6) {
                cp = c & 0x1f;
                extra = 1;
            } else if ((c >> 4) == 0xe) {
                cp = c & 0x0f;
                extra = 2;
            } else if ((c >> 3) == 0x1e) {
                cp = c & 0x07;
                extra = 3;
            } else {
                cp = '?';
                extra = 0;
            }
            i++;
            for (int k = 0; k < extra && i < len; k++, i++) {
                cp = (cp << 6) | ((unsigned char)utf8[i] & 0x3f);
            }
            if (cp >= 0x10000) {
                cp -= 0x10000;
                out.push_back((char16_t)(0xd800 + (cp >> 10)));
                out.push_back((char16_t)(0xdc00 + (cp & 0x3ff)));
            } else {
                out.push_back((char16_t)cp);
            }
        }
        return out;
    }
    static std::string toU8(const XMLCh* s, std::size_t len) {
        std::string out;
        out.reserve(len);
        for (std::size_t i = 0; i < len; i++) {
            std::uint32_t cp = s[i];
            if (cp >= 0xd800 && cp < 0xdc00 && i + 1 < len) {
                cp = 0x10000 + ((cp - 0xd800) << 10) + (s[i + 1] - 0xdc00);
                i++;
            }
            if (cp < 0x80) {
                out.push_back((char)cp);
            } else if (cp < 0x800) {
                out.push_back((char)(0xc0 | (cp >> 6)));
                out.push_back((char)(0x80 | (cp & 0x3f)));
            } else if (cp < 0x10000) {
                out.push_back((char)(0xe0 | (cp >> 12)));
                out.push_back((char)(0x80 | ((cp >> 6) & 0x3f)));
                out.push_back((char)(0x80 | (cp & 0x3f)));
            } else {
                out.push_back((char)(0xf0 | (cp >> 18)));
                out.push_back((char)(0x80 | ((cp >> 12) & 0x3f)));
                out.push_back((char)(0x80 | ((cp >> 6) & 0x3f)));
                out.push_back((char)(0x80 | (cp & 0x3f)));
            }
        }
        return out;
    }
    static XMLCh* transcode(const char* s) {
        if (s == nullptr) {
            return nullptr;
        }
        const std::u16string w = toU16(s, std::strlen(s));
        XMLCh* r = new XMLCh[w.size() + 1];
        std::memcpy(r, w.data(), w.size() * sizeof(XMLCh));
        r[w.size()] = 0;
        return r;
    }
    static XMLCh* transcode(const char* s, MemoryManager*) {
        return transcode(s);
    }
    static char* transcode(const XMLCh* s) {
        if (s == nullptr) {
            return nullptr;
        }
        const std::string n = toU8(s, stringLen(s));
        char* r = new char[n.size() + 1];
        std::memcpy(r, n.data(), n.size());
        r[n.size()] = 0;
        return r;
    }
    static void release(XMLCh** p) {
        delete[] *p;
        *p = nullptr;
    }
    static void release(char** p) {
        delete[] *p;
        *p = nullptr;
    }
};

class XMLLCPTranscoder {};
class XMLGrammarPool {
public:
    virtual ~XMLGrammarPool() {}
};
class XMLGrammarPoolImpl : public XMLGrammarPool {
public:
    XMLGrammarPoolImpl(MemoryManager* = nullptr) {}
};
class Grammar {
public:
    enum GrammarType { DTDGrammarType, SchemaGrammarType, UnKnown };
};

class XMLUni {
public:
    static inline const XMLCh fgXercesScannerName[] = u"scanner-name";
    static inline const XMLCh fgWFXMLScanner[] = u"WFXMLScanner";
    static inline const XMLCh fgIGXMLScanner[] = u"IGXMLScanner";
    static inline const XMLCh fgXercesSchema[] = u"schema";
    static inline const XMLCh fgSAX2CoreValidation[] = u"validation";
    static inline const XMLCh fgXercesDynamic[] = u"dynamic";
    static inline const XMLCh fgXercesUseCachedGrammarInParse[] = u"use-cached-grammar";
    static inline const XMLCh fgXercesHandleMultipleImports[] = u"multiple-imports";
};

// ---------------------------------------------------------------------------------------------
// exceptions
// ---------------------------------------------------------------------------------------------
class XMLException {
public:
    XMLException(const std::string& msg = "") : myMsg(XMLString::toU16(msg.data(), msg.size())) {}
    virtual ~XMLException() {}
    const XMLCh* getMessage() const {
        return myMsg.c_str();
    }
private:
    std::u16string myMsg;
};
class TranscodingException : public XMLException {};

class SAXException {
public:
    SAXException(const std::string& msg = "") : myMsg(XMLString::toU16(msg.data(), msg.size())) {}
    virtual ~SAXException() {}
    const XMLCh* getMessage() const {
        return myMsg.c_str();
    }
private:
    std::u16string myMsg;
};
class SAXParseException : public SAXException {
public:
    SAXParseException(const std::string& msg, const std::string& sysId, XMLFileLoc line, XMLFileLoc col) :
        SAXException(msg), mySysId(XMLString::toU16(sysId.data(), sysId.size())), myLine(line), myCol(col) {}
    const XMLCh* getSystemId() const {
        return mySysId.c_str();
    }
    XMLFileLoc getLineNumber() const {
        return myLine;
    }
    XMLFileLoc getColumnNumber() const {
        return myCol;
    }
private:
    std::u16string mySysId;
    XMLFileLoc myLine, myCol;
};

// ---------------------------------------------------------------------------------------------
// input sources
// ---------------------------------------------------------------------------------------------
class BinInputStream {
public:
    virtual ~BinInputStream() {}
    virtual XMLFilePos curPos() const = 0;
    virtual XMLSize_t readBytes(XMLByte* const toFill, const XMLSize_t maxToRead) = 0;
    virtual const XMLCh* getContentType() const = 0;
};

class InputSource {
public:
    InputSource(const char* systemId = "") : mySystemId(systemId == nullptr ? "" : systemId) {}
    virtual ~InputSource() {}
    virtual BinInputStream* makeStream() const = 0;
    const std::string& shimSystemId() const {
        return mySystemId;
    }
private:
    std::string mySystemId;
};

class ShimMemStream : public BinInputStream {
public:
    ShimMemStream(const XMLByte* b, XMLSize_t n) : myBuf(b), myLen(n), myPos(0) {}
    XMLFilePos curPos() const override {
        return myPos;
    }
    XMLSize_t readBytes(XMLByte* const toFill, const XMLSize_t maxToRead) override {
        const XMLSize_t n = myLen - myPos < maxToRead ? myLen - myPos : maxToRead;
        std::memcpy(toFill, myBuf + myPos, n);
        myPos += n;
        return n;
    }
    const XMLCh* getContentType() const override {
        return nullptr;
    }
private:
    const XMLByte* myBuf;
    XMLSize_t myLen, myPos;
};

class MemBufInputSource : public InputSource {
public:
    MemBufInputSource(const XMLByte* b, XMLSize_t n, const char* id, bool = false) : InputSource(id), myBuf(b), myLen(n) {}
    BinInputStream* makeStream() const override {
        return new ShimMemStream(myBuf, myLen);
    }
private:
    const XMLByte* myBuf;
    XMLSize_t myLen;
};

class ShimFileStream : public BinInputStream {
public:
    ShimFileStream(const std::string& path) : myIn(path.c_str(), std::ios::in | std::ios::binary) {}
    bool good() const {
        return myIn.good();
    }
    XMLFilePos curPos() const override {
        return (XMLFilePos)const_cast<std::ifstream&>(myIn).tellg();
    }
    XMLSize_t readBytes(XMLByte* const toFill, const XMLSize_t maxToRead) override {
        myIn.read((char*)toFill, (std::streamsize)maxToRead);
        return (XMLSize_t)myIn.gcount();
    }
    const XMLCh* getContentType() const override {
        return nullptr;
    }
private:
    std::ifstream myIn;
};

class LocalFileInputSource : public InputSource {
public:
    LocalFileInputSource(const XMLCh* path) : InputSource(XMLString::toU8(path, XMLString::stringLen(path)).c_str()) {}
    BinInputStream* makeStream() const override {
        return new ShimFileStream(shimSystemId());
    }
};

class EntityResolver {
public:
    virtual ~EntityResolver() {}
    virtual InputSource* resolveEntity(const XMLCh* const publicId, const XMLCh* const systemId) = 0;
};

class XMLPScanToken {};

// ---------------------------------------------------------------------------------------------
// attributes (one object serves both the SAX1 AttributeList and the SAX2 Attributes interface)
// ---------------------------------------------------------------------------------------------
class ShimAttrStore {
public:
    std::vector<std::u16string> names, values;
    int find(const XMLCh* name) const {
        for (std::size_t i = 0; i < names.size(); i++) {
            if (XMLString::equals(names[i].c_str(), name)) {
                return (int)i;
            }
        }
        return -1;
    }
};

class Attributes {
public:
    virtual ~Attributes() {}
    XMLSize_t getLength() const {
        return myStore.names.size();
    }
    const XMLCh* getLocalName(const XMLSize_t i) const {
        return i < myStore.names.size() ? myStore.names[i].c_str() : nullptr;
    }
    const XMLCh* getQName(const XMLSize_t i) const {
        return getLocalName(i);
    }
    const XMLCh* getValue(const XMLSize_t i) const {
        return i < myStore.values.size() ? myStore.values[i].c_str() : nullptr;
    }
    const XMLCh* getValue(const XMLCh* const qname) const {
        const int i = myStore.find(qname);
        return i < 0 ? nullptr : myStore.values[i].c_str();
    }
    int getIndex(const XMLCh* const qname) const {
        return myStore.find(qname);
    }
    ShimAttrStore myStore;
};

class AttributeList {
public:
    virtual ~AttributeList() {}
    XMLSize_t getLength() const {
        return myStore.names.size();
    }
    const XMLCh* getName(const XMLSize_t i) const {
        return i < myStore.names.size() ? myStore.names[i].c_str() : nullptr;
    }
    const XMLCh* getValue(const XMLSize_t i) const {
        return i < myStore.values.size() ? myStore.values[i].c_str() : nullptr;
    }
    const XMLCh* getValue(const XMLCh* const name) const {
        const int i = myStore.find(name);
        return i < 0 ? nullptr : myStore.values[i].c_str();
    }
    const XMLCh* getValue(const char* const name) const {
        const std::u16string w = XMLString::toU16(name, std::strlen(name));
        return getValue(w.c_str());
    }
    ShimAttrStore myStore;
};

// ---------------------------------------------------------------------------------------------
// handler interfaces
// ---------------------------------------------------------------------------------------------
class ErrorHandler {
public:
    virtual ~ErrorHandler() {}
    virtual void warning(const SAXParseException&) {}
    virtual void error(const SAXParseException&) {}
    virtual void fatalError(const SAXParseException& e) {
        throw e;
    }
    virtual void resetErrors() {}
};

class ContentHandler {
public:
    virtual ~ContentHandler() {}
    virtual void startElement(const XMLCh* const, const XMLCh* const, const XMLCh* const, const Attributes&) {}
    virtual void endElement(const XMLCh* const, const XMLCh* const, const XMLCh* const) {}
    virtual void characters(const XMLCh* const, const XMLSize_t) {}
    virtual void startDocument() {}
    virtual void endDocument() {}
};

class DefaultHandler : public ContentHandler, public ErrorHandler, public EntityResolver {
public:
    InputSource* resolveEntity(const XMLCh* const, const XMLCh* const) override {
        return nullptr;
    }
};

class DocumentHandler {
public:
    virtual ~DocumentHandler() {}
    virtual void startElement(const XMLCh* const, AttributeList&) {}
    virtual void endElement(const XMLCh* const) {}
    virtual void characters(const XMLCh* const, const XMLSize_t) {}
    virtual void startDocument() {}
    virtual void endDocument() {}
};

class HandlerBase : public DocumentHandler, public ErrorHandler, public EntityResolver {
public:
    InputSource* resolveEntity(const XMLCh* const, const XMLCh* const) override {
        return nullptr;
    }
};

// ---------------------------------------------------------------------------------------------
// the expat pump: turns an input stream into a queue of events
// ---------------------------------------------------------------------------------------------
struct ShimEvent {
    enum Kind { START, END, CHARS, FATAL } kind;
    std::u16string name;       // element name, character data or error text
    ShimAttrStore attrs;
    XMLFileLoc line = 0, col = 0;
};

class ShimPump {
public:
    ShimPump(BinInputStream* in, const std::string& sysId) : myIn(in), mySysId(sysId), myDone(false) {
        myParser = XML_ParserCreate(nullptr);
        XML_SetUserData(myParser, this);
        XML_SetElementHandler(myParser, &ShimPump::onStart, &ShimPump::onEnd);
        XML_SetCharacterDataHandler(myParser, &ShimPump::onChars);
    }
    ~ShimPump() {
        XML_ParserFree(myParser);
    }
    ShimPump(const ShimPump&) = delete;
    ShimPump& operator=(const ShimPump&) = delete;

    /// @brief fetches the next event; returns false at the end of the document
    bool next(ShimEvent& ev) {
        while (myQueue.empty()) {
            if (myDone) {
                return false;
            }
            XMLByte buf[1 << 16];
            const XMLSize_t n = myIn->readBytes(buf, sizeof(buf));
            const bool final = n == 0;
            if (XML_Parse(myParser, (const char*)buf, (int)n, final ? 1 : 0) == XML_STATUS_ERROR) {
                ShimEvent e;
                e.kind = ShimEvent::FATAL;
                const char* msg = XML_ErrorString(XML_GetErrorCode(myParser));
                e.name = XMLString::toU16(msg, std::strlen(msg));
                e.line = XML_GetCurrentLineNumber(myParser) - 1;
                e.col = XML_GetCurrentColumnNumber(myParser);
                myQueue.push_back(std::move(e));
                myDone = true;
            } else if (final) {
                myDone = true;
            }
        }
        ev = std::move(myQueue.front());
        myQueue.pop_front();
        return true;
    }
    const std::string& sysId() const {
        return mySysId;
    }

private:
    static void XMLCALL onStart(void* ud, const XML_Char* name, const XML_Char** atts) {
        ShimPump* self = (ShimPump*)ud;
        self->flushChars();
        ShimEvent e;
        e.kind = ShimEvent::START;
        e.name = XMLString::toU16(name, std::strlen(name));
        for (int i = 0; atts[i] != nullptr; i += 2) {
            e.attrs.names.push_back(XMLString::toU16(atts[i], std::strlen(atts[i])));
            e.attrs.values.push_back(XMLString::toU16(atts[i + 1], std::strlen(atts[i + 1])));
        }
        self->myQueue.push_back(std::move(e));
    }
    static void XMLCALL onEnd(void* ud, const XML_Char* name) {
        ShimPump* self = (ShimPump*)ud;
        self->flushChars();
        ShimEvent e;
        e.kind = ShimEvent::END;
        e.name = XMLString::toU16(name, std::strlen(name));
        self->myQueue.push_back(std::move(e));
    }
    static void XMLCALL onChars(void* ud, const XML_Char* s, int len) {
        ((ShimPump*)ud)->myChars.append(s, (std::size_t)len);
    }
    void flushChars() {
        if (!myChars.empty()) {
            ShimEvent e;
            e.kind = ShimEvent::CHARS;
            e.name = XMLString::toU16(myChars.data(), myChars.size());
            myQueue.push_back(std::move(e));
            myChars.clear();
        }
    }

    XML_Parser myParser;
    std::unique_ptr<BinInputStream> myIn;
    std::string mySysId;
    std::string myChars;
    std::deque<ShimEvent> myQueue;
    bool myDone;
};

inline ShimPump* shimOpenFile(const char* path) {
    ShimFileStream* fs = new ShimFileStream(path);
    if (!fs->good()) {
        delete fs;
        throw SAXParseException(std::string("unable to open primary document entity '") + path + "'", path, 0, 0);
    }
    return new ShimPump(fs, path);
}

// ---------------------------------------------------------------------------------------------
// SAX2 reader
// ---------------------------------------------------------------------------------------------
class SAX2XMLReader {
public:
    SAX2XMLReader() : myContent(nullptr), myError(nullptr) {}
    virtual ~SAX2XMLReader() {}
    void setContentHandler(ContentHandler* h) {
        myContent = h;
    }
    void setErrorHandler(ErrorHandler* h) {
        myError = h;
    }
    void setEntityResolver(EntityResolver*) {}
    void setProperty(const XMLCh*, void*) {}
    void setFeature(const XMLCh*, bool) {}
    bool loadGrammar(const char*, Grammar::GrammarType, bool) {
        return true;
    }
    void parse(const char* path) {
        myPump.reset(shimOpenFile(path));
        drain();
    }
    void parse(const InputSource& src) {
        myPump.reset(new ShimPump(src.makeStream(), src.shimSystemId()));
        drain();
    }
    bool parseFirst(const char* path, XMLPScanToken&) {
        myPump.reset(shimOpenFile(path));
        return true;
    }
    bool parseFirst(const InputSource& src, XMLPScanToken&) {
        myPump.reset(new ShimPump(src.makeStream(), src.shimSystemId()));
        return true;
    }
    bool parseNext(XMLPScanToken&) {
        return myPump != nullptr && step();
    }
    void parseReset(XMLPScanToken&) {
        myPump.reset();
    }

private:
    void drain() {
        // the handler may start a nested parse on this very reader object only through another
        // SUMOSAXReader (XMLSubSys keeps a reader stack), so holding the pump locally is safe
        std::unique_ptr<ShimPump> pump = std::move(myPump);
        ShimEvent ev;
        while (pump->next(ev)) {
            dispatch(ev, *pump);
        }
    }
    bool step() {
        ShimEvent ev;
        if (!myPump->next(ev)) {
            return false;
        }
        dispatch(ev, *myPump);
        return true;
    }
    void dispatch(ShimEvent& ev, ShimPump& pump) {
        static const XMLCh empty[] = u"";
        switch (ev.kind) {
            case ShimEvent::START:
                myAttrs.myStore = std::move(ev.attrs);
                if (myContent != nullptr) {
                    myContent->startElement(empty, ev.name.c_str(), ev.name.c_str(), myAttrs);
                }
                break;
            case ShimEvent::END:
                if (myContent != nullptr) {
                    myContent->endElement(empty, ev.name.c_str(), ev.name.c_str());
                }
                break;
            case ShimEvent::CHARS:
                if (myContent != nullptr) {
                    myContent->characters(ev.name.c_str(), ev.name.size());
                }
                break;
            case ShimEvent::FATAL: {
                const SAXParseException e(XMLString::toU8(ev.name.data(), ev.name.size()), pump.sysId(), ev.line, ev.col);
                if (myError != nullptr) {
                    myError->fatalError(e);
                } else {
                    throw e;
                }
                break;
            }
        }
    }
    ContentHandler* myContent;
    ErrorHandler* myError;
    std::unique_ptr<ShimPump> myPump;
    Attributes myAttrs;   // persists until the next start tag, like the Xerces reader's own vector
};

class XMLReaderFactory {
public:
    static SAX2XMLReader* createXMLReader(MemoryManager* = nullptr, XMLGrammarPool* = nullptr) {
        return new SAX2XMLReader();
    }
};

// ---------------------------------------------------------------------------------------------
// SAX1 parser
// ---------------------------------------------------------------------------------------------
class SAXParser {
public:
    enum ValSchemes { Val_Never, Val_Always, Val_Auto };
    SAXParser() : myDoc(nullptr), myError(nullptr) {}
    void setValidationScheme(ValSchemes) {}
    void setDisableDefaultEntityResolution(bool) {}
    void setDocumentHandler(DocumentHandler* h) {
        myDoc = h;
    }
    void setErrorHandler(ErrorHandler* h) {
        myError = h;
    }
    void parse(const char* path) {
        myPump.reset(shimOpenFile(path));
        while (step()) {}
    }
    void parse(const InputSource& src) {
        myPump.reset(new ShimPump(src.makeStream(), src.shimSystemId()));
        while (step()) {}
    }
    bool parseFirst(const char* path, XMLPScanToken&) {
        myPump.reset(shimOpenFile(path));
        return true;
    }
    bool parseFirst(const InputSource& src, XMLPScanToken&) {
        myPump.reset(new ShimPump(src.makeStream(), src.shimSystemId()));
        return true;
    }
    bool parseNext(XMLPScanToken&) {
        return myPump != nullptr && step();
    }

private:
    bool step() {
        ShimEvent ev;
        if (!myPump->next(ev)) {
            return false;
        }
        switch (ev.kind) {
            case ShimEvent::START:
                myAttrs.myStore = std::move(ev.attrs);
                if (myDoc != nullptr) {
                    myDoc->startElement(ev.name.c_str(), myAttrs);
                }
                break;
            case ShimEvent::END:
                if (myDoc != nullptr) {
                    myDoc->endElement(ev.name.c_str());
                }
                break;
            case ShimEvent::CHARS:
                if (myDoc != nullptr) {
                    myDoc->characters(ev.name.c_str(), ev.name.size());
                }
                break;
            case ShimEvent::FATAL: {
                const SAXParseException e(XMLString::toU8(ev.name.data(), ev.name.size()), myPump->sysId(), ev.line, ev.col);
                if (myError != nullptr) {
                    myError->fatalError(e);
                } else {
                    throw e;
                }
                break;
            }
        }
        return true;
    }
    DocumentHandler* myDoc;
    ErrorHandler* myError;
    std::unique_ptr<ShimPump> myPump;
    AttributeList myAttrs;
};

} // namespace xercesc
