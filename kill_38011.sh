ssh test@127.0.0.1 "kill -9 9555"
ssh test@127.0.0.1 "kill -9 9556"
ssh test@127.0.0.1 "kill -9 9557"
kill -9 8873
