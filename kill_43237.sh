ssh test@127.0.0.1 "kill -9 9369"
kill -9 8873
