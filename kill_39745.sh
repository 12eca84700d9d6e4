ssh test@127.0.0.1 "kill -9 5070"
ssh test@127.0.0.1 "kill -9 5071"
kill -9 5044
